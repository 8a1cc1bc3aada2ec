// tcgen05 policy path (TFODE_POLICY_TC): TF_model<float>::infer (tf_cpp_binding.cc:70-113) for the actor of
// scripts/rover_training_1.py:26-35 (17 -> 512 relu -> 512 relu -> 2 tanh) on the 5th-generation tensor cores.
//
// One CTA = 128 environments (the M of a 128xNx16 UMMA tile = the 128 TMEM lanes).  Per CTA:
//   gather  : thread r gathers its environment's 17 scaled observations (parked list or direct rows), rounds them to
//             bf16 and stores them as the A operand (K padded to 32) in the UMMA canonical K-major no-swizzle layout
//   layer 1 : W1 (bf16, canonical, 32 KB) arrives by one bulk TMA copy; 4 x tcgen05.mma (N=256 halves x 2 K-steps)
//             accumulate H1[128][512] in fp32 into all 512 TMEM columns
//   epilogue: every thread reads its lane's 512 columns with tcgen05.ld, adds b1, ReLU, rounds to bf16 and writes the
//             layer-2 A operand (128 KB) straight back into shared memory in the canonical layout
//   layer 2 : W2 (512 KB) streams through a 2-stage 32 KB ring of bulk TMA copies tracked by mbarriers; 64 x
//             tcgen05.mma into the same TMEM columns; tcgen05.commit releases ring slots and signals the epilogue
//   epilogue: b2 + ReLU in fp32, the 512 -> 2 output layer as two fp32 dot products per thread, tanh, and the two
//             actions go straight into the environment record (policy_act, parked = 2) that the step kernel turns
//             into the steering-rate / boggie-torque setpoints (src/rover_1_tf.cc:184-185).
// Operand layout (bf16, "core matrix" = 8 rows x 16 bytes, stored as 128 contiguous bytes):
//   element (row, k) of an operand with R rows lives at  (k/8) * (R/8)*128 + (row/8)*128 + (row%8)*16 + (k%8)*2
//   => stride-dimension byte offset (next 8 rows) = 128, leading-dimension byte offset (next 8 k) = R*16.
#pragma once
#include <cuda_bf16.h>

#include <vector>

#include "tfode_device.cuh"

namespace tfode {

constexpr int TC_ROWS = 128;                   // environments per CTA
constexpr int TC_H = TFODE_MLP_HIDDEN;         // 512
constexpr int TC_K1 = 32;                      // 17 padded to two K=16 steps
constexpr int TC_SLAB_K = 32;                  // K extent of one W2 ring slot
constexpr int TC_SLAB_BYTES = TC_H * TC_SLAB_K * 2;   // 32 KB
constexpr int TC_NSLAB = TC_H / TC_SLAB_K;     // 16
constexpr int TC_A2_BYTES = TC_ROWS * TC_H * 2;       // 128 KB
constexpr int TC_A1_BYTES = TC_ROWS * TC_K1 * 2;      // 8 KB
// shared memory map (bytes)
constexpr int TC_OFF_A2 = 0;
constexpr int TC_OFF_RING = TC_OFF_A2 + TC_A2_BYTES;            // 2 slots; slot 0 first holds W1
constexpr int TC_OFF_A1 = TC_OFF_RING + 2 * TC_SLAB_BYTES;
constexpr int TC_OFF_B1 = TC_OFF_A1 + TC_A1_BYTES;              // fp32 b1[512]
constexpr int TC_OFF_B2 = TC_OFF_B1 + TC_H * 4;                 // fp32 b2[512]
constexpr int TC_OFF_W3 = TC_OFF_B2 + TC_H * 4;                 // fp32 W3[512][2]
constexpr int TC_OFF_BAR = TC_OFF_W3 + TC_H * 2 * 4;            // mbarriers: full[2], empty[2], done
constexpr int TC_OFF_TMEM = TC_OFF_BAR + 8 * 8;
constexpr int TC_SMEM_BYTES = TC_OFF_TMEM + 16;

struct PolicyTC {
  __nv_bfloat16* w1c = nullptr;  // [4 k-cores][64 n-cores][8][8]
  __nv_bfloat16* w2c = nullptr;  // [64 k-cores][64 n-cores][8][8]
  bool attr_set = false;
};

namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
// 1-D bulk TMA copy global -> shared, completion counted in bytes on an mbarrier
__device__ __forceinline__ void tma_bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// K-major, no swizzle: start>>4 | LBO>>4 @16 | SBO>>4 @32 | version 1 @46
__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((addr & 0x3FFFF) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) | (1ull << 46);
}
// kind::f16 instruction descriptor: D fp32, A/B bf16, both K-major, M = 128, N = 256
constexpr uint32_t IDESC_128x256 = (1u << 4) | (1u << 7) | (1u << 10) | ((256u >> 3) << 17) | ((128u >> 4) << 24);

__device__ __forceinline__ void mma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void mma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// 32 lanes x 32 consecutive fp32 columns -> 32 registers per thread (thread = lane)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
        "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
        "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);  // .x = lo (low half), .y = hi
  return *reinterpret_cast<uint32_t*>(&t);
}

}  // namespace tc

// in_all: [*,17] scaled observations; list/n_list: parked environments (device) or nullptr for direct rows [0, m_direct)
__global__ void __launch_bounds__(TC_ROWS, 1)
policy_tc_kernel(const float* __restrict__ in_all, const int* __restrict__ list, const int* __restrict__ n_list, int m_direct,
                 const __nv_bfloat16* __restrict__ w1c, const __nv_bfloat16* __restrict__ w2c, const float* __restrict__ b1,
                 const float* __restrict__ b2, const float* __restrict__ W3, const float* __restrict__ b3, EnvRec* rec, float* act_direct) {
  using namespace tc;
  extern __shared__ __align__(1024) uint8_t smem[];
  const int m = list ? *n_list : m_direct;
  const int row0 = blockIdx.x * TC_ROWS;
  if (row0 >= m) return;                       // uniform per CTA, before any allocation
  const int t = threadIdx.x, warp = t >> 5;
  const int row = row0 + t;
  const bool valid = row < m;
  const int src = valid ? (list ? list[row] : row) : -1;

  const uint32_t sA2 = smem_u32(smem + TC_OFF_A2), sRing = smem_u32(smem + TC_OFF_RING), sA1 = smem_u32(smem + TC_OFF_A1);
  const uint32_t barFull0 = smem_u32(smem + TC_OFF_BAR), barEmpty0 = barFull0 + 16, barDone = barFull0 + 32;
  float* sb1 = reinterpret_cast<float*>(smem + TC_OFF_B1);
  float* sb2 = reinterpret_cast<float*>(smem + TC_OFF_B2);
  float* sW3 = reinterpret_cast<float*>(smem + TC_OFF_W3);
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + TC_OFF_TMEM);

  if (t == 0) {
    mbar_init(barFull0, 1); mbar_init(barFull0 + 8, 1);
    mbar_init(barEmpty0, 1); mbar_init(barEmpty0 + 8, 1);
    mbar_init(barDone, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    // W1 into ring slot 0 while the rows are gathered
    mbar_expect_tx(barFull0, TC_SLAB_BYTES);
    tma_bulk_g2s(sRing, w1c, TC_SLAB_BYTES, barFull0);
  }
  if (warp == 1) {  // one warp owns the TMEM allocation (all 512 columns: 128 lanes x 512 fp32 accumulators)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem + TC_OFF_TMEM)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // biases and the output layer, fp32
  for (int i = t; i < TC_H; i += TC_ROWS) { sb1[i] = b1[i]; sb2[i] = b2[i]; sW3[2 * i] = W3[2 * i]; sW3[2 * i + 1] = W3[2 * i + 1]; }
  // A1: this thread's row, K = 32 (17 observations, zero padded), 4 k-cores of 16 bytes
  {
    float x[TC_K1];
#pragma unroll
    for (int k = 0; k < TC_K1; k++) x[k] = (valid && k < 17) ? in_all[(size_t)src * 17 + k] : 0.0f;
#pragma unroll
    for (int kc = 0; kc < TC_K1 / 8; kc++) {
      uint4 q = make_uint4(pack_bf16(x[kc * 8], x[kc * 8 + 1]), pack_bf16(x[kc * 8 + 2], x[kc * 8 + 3]), pack_bf16(x[kc * 8 + 4], x[kc * 8 + 5]),
                           pack_bf16(x[kc * 8 + 6], x[kc * 8 + 7]));
      *reinterpret_cast<uint4*>(smem + TC_OFF_A1 + kc * (TC_ROWS * 16) + t * 16) = q;
    }
  }
  fence_proxy_async();     // generic-proxy stores of A1 must be visible to the tensor core (async proxy)
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  // ---- layer 1: H1 = X(128x32) . W1^T, N in two halves of 256 columns
  if (t == 0) {
    mbar_wait(barFull0, 0);
    tc_fence_after();
#pragma unroll
    for (int ks = 0; ks < TC_K1 / 16; ks++)
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const uint64_t ad = make_desc(sA1 + ks * 2 * (TC_ROWS * 16), TC_ROWS * 16, 128);
        const uint64_t bd = make_desc(sRing + ks * 2 * (TC_H * 16) + h * (256 / 8) * 128, TC_H * 16, 128);
        mma_bf16(tmem + h * 256, ad, bd, IDESC_128x256, ks > 0);
      }
    mma_commit(barDone);
  }
  mbar_wait(barDone, 0);
  tc_fence_after();
  if (t == 0) {  // ring slot 0 (W1) is free again: start streaming W2 slabs 0 and 1 under the epilogue
    mbar_expect_tx(barFull0, TC_SLAB_BYTES);
    tma_bulk_g2s(sRing, w2c, TC_SLAB_BYTES, barFull0);
    mbar_expect_tx(barFull0 + 8, TC_SLAB_BYTES);
    tma_bulk_g2s(sRing + TC_SLAB_BYTES, reinterpret_cast<const uint8_t*>(w2c) + TC_SLAB_BYTES, TC_SLAB_BYTES, barFull0 + 8);
  }
  // ---- epilogue 1: bias + ReLU + bf16, written as the layer-2 A operand
  const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
  for (int c = 0; c < TC_H / 32; c++) {
    uint32_t v[32];
    tmem_ld32(lane_base + c * 32, v);
#pragma unroll
    for (int g = 0; g < 4; g++) {
      float f[8];
#pragma unroll
      for (int i = 0; i < 8; i++) { float a = __uint_as_float(v[g * 8 + i]) + sb1[c * 32 + g * 8 + i]; f[i] = a > 0.0f ? a : 0.0f; }
      uint4 q = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
      *reinterpret_cast<uint4*>(smem + TC_OFF_A2 + (c * 4 + g) * (TC_ROWS * 16) + t * 16) = q;
    }
  }
  fence_proxy_async();
  tc_fence_before();       // the tcgen05.ld reads of H1 are ordered before the MMAs that overwrite those columns
  __syncthreads();
  tc_fence_after();

  // ---- layer 2: H2 = relu(H1)(128x512) . W2^T, W2 streamed in 16 slabs of K = 32 through the 2-slot ring
  if (t == 0) {
#pragma unroll 1
    for (int s = 0; s < TC_NSLAB; s++) {
      const int slot = s & 1;
      const uint32_t ph = (uint32_t)(s >> 1) & 1u;
      // full[slot] completed once already for slot 0 (W1): its phase for W2 slab s is shifted by one
      mbar_wait(barFull0 + 8 * slot, slot == 0 ? (ph ^ 1u) : ph);
      tc_fence_after();
#pragma unroll
      for (int ks = 0; ks < TC_SLAB_K / 16; ks++)
#pragma unroll
        for (int h = 0; h < 2; h++) {
          const uint64_t ad = make_desc(sA2 + (s * (TC_SLAB_K / 8) + ks * 2) * (TC_ROWS * 16), TC_ROWS * 16, 128);
          const uint64_t bd = make_desc(sRing + slot * TC_SLAB_BYTES + ks * 2 * (TC_H * 16) + h * (256 / 8) * 128, TC_H * 16, 128);
          mma_bf16(tmem + h * 256, ad, bd, IDESC_128x256, (s | ks) > 0);
        }
      mma_commit(barEmpty0 + 8 * slot);       // arrives when the MMAs above have finished reading the slot
      // refill the OTHER slot (slab s-1's) with slab s+1 while slab s is in the tensor pipe
      if (s >= 1 && s + 1 < TC_NSLAB) {
        const int oslot = slot ^ 1;
        mbar_wait(barEmpty0 + 8 * oslot, (uint32_t)((s - 1) >> 1) & 1u);
        mbar_expect_tx(barFull0 + 8 * oslot, TC_SLAB_BYTES);
        tma_bulk_g2s(sRing + oslot * TC_SLAB_BYTES, reinterpret_cast<const uint8_t*>(w2c) + (size_t)(s + 1) * TC_SLAB_BYTES, TC_SLAB_BYTES, barFull0 + 8 * oslot);
      }
    }
    mma_commit(barDone);
  }
  mbar_wait(barDone, 1);
  tc_fence_after();

  // ---- epilogue 2: bias + ReLU, output layer 512 -> 2 in fp32, tanh, actions into the environment record
  float s0 = 0.0f, s1 = 0.0f;
#pragma unroll 1
  for (int c = 0; c < TC_H / 32; c++) {
    uint32_t v[32];
    tmem_ld32(lane_base + c * 32, v);
#pragma unroll
    for (int i = 0; i < 32; i++) {
      float a = __uint_as_float(v[i]) + sb2[c * 32 + i];
      a = a > 0.0f ? a : 0.0f;
      s0 = fmaf(a, sW3[2 * (c * 32 + i)], s0);
      s1 = fmaf(a, sW3[2 * (c * 32 + i) + 1], s1);
    }
  }
  if (valid) {
    const float a0 = tfn::tanh(s0 + b3[0]), a1 = tfn::tanh(s1 + b3[1]);
    if (list) { EnvRec& e = rec[src]; e.c.policy_act[0] = a0; e.c.policy_act[1] = a1; e.c.parked = 2; }
    else { act_direct[(size_t)row * 2] = a0; act_direct[(size_t)row * 2 + 1] = a1; }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
}

// ---- host side: operand preparation and launch ----
inline uint16_t tc_bf16_rn(float f) {  // round-to-nearest-even, same as __float2bfloat16_rn for finite inputs
  uint32_t u; memcpy(&u, &f, 4);
  if ((u & 0x7F800000u) == 0x7F800000u) return (uint16_t)((u >> 16) | ((u & 0xFFFFu) ? 0x40u : 0u));
  u += 0x7FFFu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
// W is the Keras kernel [K_in][512] (tf_cpp_binding weights, scripts/rover_training_1.py:26-35); the B operand is W^T
// ([n][k], K-major) in the canonical layout described at the top, K padded to kpad.
inline void tc_pack_operand(const float* W, int k_in, int kpad, std::vector<uint16_t>& out) {
  out.assign((size_t)kpad * TC_H, 0);
  for (int k = 0; k < k_in; k++)
    for (int n = 0; n < TC_H; n++)
      out[(size_t)(k / 8) * (TC_H / 8) * 64 + (size_t)(n / 8) * 64 + (n % 8) * 8 + (k % 8)] = tc_bf16_rn(W[(size_t)k * TC_H + n]);
}
inline int policy_tc_upload(PolicyTC& tcw, const float* w_host, cudaStream_t stream) {
  const size_t H = TC_H;
  const float* W1 = w_host; const float* W2 = w_host + 17 * H + H;
  std::vector<uint16_t> p1, p2;
  tc_pack_operand(W1, 17, TC_K1, p1);
  tc_pack_operand(W2, TC_H, TC_H, p2);
  if (!tcw.w1c && cudaMalloc(&tcw.w1c, p1.size() * 2) != cudaSuccess) return -1;
  if (!tcw.w2c && cudaMalloc(&tcw.w2c, p2.size() * 2) != cudaSuccess) return -1;
  if (cudaMemcpyAsync(tcw.w1c, p1.data(), p1.size() * 2, cudaMemcpyHostToDevice, stream) != cudaSuccess) return -1;
  if (cudaMemcpyAsync(tcw.w2c, p2.data(), p2.size() * 2, cudaMemcpyHostToDevice, stream) != cudaSuccess) return -1;
  if (cudaStreamSynchronize(stream) != cudaSuccess) return -1;   // p1/p2 are pageable and die with this scope
  return 0;
}
inline void policy_tc_free(PolicyTC& tcw) { cudaFree(tcw.w1c); cudaFree(tcw.w2c); tcw.w1c = tcw.w2c = nullptr; }
// returns the CUDA status of the launch (cudaErrorInvalidValue when no weights were uploaded)
inline cudaError_t policy_tc_launch(PolicyTC& tcw, const float* in, const int* list, const int* n_list, int m_direct, int max_rows, const float* b1, const float* b2,
                            const float* W3, const float* b3, EnvRec* rec, float* act_direct, cudaStream_t stream) {
  if (!tcw.w1c || !tcw.w2c) return cudaErrorInvalidValue;
  if (!tcw.attr_set) {
    const cudaError_t e = cudaFuncSetAttribute(policy_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES);
    if (e != cudaSuccess) return e;
    tcw.attr_set = true;
  }
  const int blocks = (max_rows + TC_ROWS - 1) / TC_ROWS;
  policy_tc_kernel<<<blocks, TC_ROWS, TC_SMEM_BYTES, stream>>>(in, list, n_list, m_direct, tcw.w1c, tcw.w2c, b1, b2, W3, b3, rec, act_direct);
  return cudaGetLastError();
}
}  // namespace tfode
