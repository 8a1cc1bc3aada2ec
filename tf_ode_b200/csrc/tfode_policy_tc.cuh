// tcgen05 policy path (TFODE_POLICY_TC): TF_model<float>::infer (tf_cpp_binding.cc:70-113) for the actor of
// scripts/rover_training_1.py:26-35 (17 -> 512 relu -> 512 relu -> 2 tanh) on the 5th-generation tensor cores.
//
// One CTA = 128 environments (the M of a 128xNx16 UMMA tile = the 128 TMEM lanes), ten warps in three roles:
//   warp 8  producer: W1 (32 KB, one bulk TMA copy) and then W2 as 32 slabs of 16 KB (K = 32 x N = 256, one N half after the other) through a
//           5-slot ring tracked by full/empty mbarriers -- it starts at kernel entry, so the stream from L2 runs under the gather, layer 1 and
//           its epilogue and never waits for anything but a free slot.
//   warp 9  MMA issuer (and TMEM owner): 4 x tcgen05.mma for layer 1 (all 512 TMEM columns), then 2 per slab for layer 2, all from ONE elected
//           lane running the whole (unrolled) loop; tcgen05.commit frees ring slots and signals "N half 0 done" / "N half 1 done".
//           (Issued from an `if (lane == 0)` region instead, every tcgen05 / TMA instruction is wrapped by the compiler in a waterfall loop
//           of R2UR moves, and the tensor pipe idles half of the time.)
//   warps 0-7  gather + epilogues; warp w reads TMEM lanes 32 (w % 4) .., the two warps of a lane quarter split the columns
//     gather    : thread r < 128 rounds its environment's 17 scaled observations (parked list or direct rows) to bf16 and stores them as the
//                 layer-1 A operand (K padded to 32) in the UMMA canonical K-major no-swizzle layout; k = 17, 18 hold 1.0 and meet the two
//                 bf16 pieces (hi + lo) of b1 that the host packed into W1's padding rows, so layer 1 needs no bias pass
//     epilogue 1: tcgen05.ld (two 32-column chunks in flight), ReLU + bf16 in one cvt.rn.relu.bf16x2.f32 per pair, written straight back to
//                 shared memory as the layer-2 A operand (128 KB, canonical layout; W1 and the layer-1 A operand lived in the same bytes);
//                 the drained TMEM columns are refilled with b2 (tcgen05.st), so layer 2 accumulates on top of its bias
//     epilogue 2: N half 0 (columns 0-255) is reduced while the tensor core works on N half 1: ReLU in fp32, the 512 -> 2 output layer as
//                 packed fma.rn.f32x2 dot products per thread; then half 1, tanh, and the two actions go straight into the environment
//                 record (policy_act, parked = 2) that the step kernel turns into the steering-rate / boggie-torque setpoints
//                 (src/rover_1_tf.cc:184-185).
// Where the time goes (tools/policy_tc_timeline.cu, 16384 rows, clk per CTA of 14800): prologue to layer-1 done 2850, first half of epilogue 1
// 1350, layer 2 8450 = 132 clk per 128x256x16 MMA, the tensor pipe's own rate (the second half of epilogue 1 and epilogue 2 of N half 0 run
// under it), tail 2100.  Every CTA asks L2 for the same W2 slab at about the same time, and that is what L2 serves best: with the K order
// rotated per CTA (tried against slice hot-spotting) the 128 streams ran into the L2 slices' ~6300 B/clk and layer 2 took 162 clk per MMA.
// 10.0 us per launch = 890 TFLOP/s, 63 % of the sustained (53 % of the burst) bf16 rate of this GPU; 1086 TFLOP/s at 131072 rows
// (profiles/r02_policy_bench.txt, r02_ncu_policy_tc_kernel.md).
// Operand layout (bf16, "core matrix" = 8 rows x 16 bytes, stored as 128 contiguous bytes):
//   element (row, k) of an operand with R rows lives at  (k/8) * (R/8)*128 + (row/8)*128 + (row%8)*16 + (k%8)*2
//   => stride-dimension byte offset (next 8 rows) = 128, leading-dimension byte offset (next 8 k) = R*16.
// W2 in global memory is stored slab by slab ([N half][K slab of 32] -> a canonical operand with R = 256 rows), so a slab is 16 KB contiguous.
#pragma once
#include <cuda_bf16.h>

#include <vector>

#include "tfode_device.cuh"

namespace tfode {

constexpr int TC_ROWS = 128;                   // environments per CTA
constexpr int TC_THREADS = 320;                // 8 epilogue warps + producer + MMA issuer
constexpr int TC_H = TFODE_MLP_HIDDEN;         // 512
constexpr int TC_K1 = 32;                      // 17 padded to two K=16 steps
constexpr int TC_SLAB_K = 32;                  // K extent of one W2 ring slot
constexpr int TC_SLAB_N = 256;                 // N extent of one W2 ring slot (one MMA's N)
constexpr int TC_SLAB_BYTES = TC_SLAB_N * TC_SLAB_K * 2;      // 16 KB
constexpr int TC_NSLAB = (TC_H / TC_SLAB_K) * (TC_H / TC_SLAB_N);   // 32: slabs 0-15 are N half 0, 16-31 N half 1
#ifndef TFODE_TC_NSLOT
#define TFODE_TC_NSLOT 5
#endif
constexpr int TC_NSLOT = TFODE_TC_NSLOT;
constexpr int TC_W1_BYTES = TC_H * TC_K1 * 2;         // 32 KB
constexpr int TC_A2_BYTES = TC_ROWS * TC_H * 2;       // 128 KB
constexpr int TC_A1_BYTES = TC_ROWS * TC_K1 * 2;      // 8 KB
// shared memory map (bytes)
constexpr int TC_OFF_A2 = 0;                                    // layer-2 A operand; before epilogue 1: W1 at +0, layer-1 A operand at +32 KB
constexpr int TC_OFF_W1 = TC_OFF_A2;
constexpr int TC_OFF_A1 = TC_OFF_W1 + TC_W1_BYTES;
constexpr int TC_OFF_RING = TC_OFF_A2 + TC_A2_BYTES;            // 5 slots of 16 KB
constexpr int TC_OFF_B2 = TC_OFF_RING + TC_NSLOT * TC_SLAB_BYTES;   // fp32 b2[512]
constexpr int TC_OFF_W3 = TC_OFF_B2 + TC_H * 4;                 // fp32 W3 transposed: [2][512]
constexpr int TC_OFF_PART = TC_OFF_W3 + TC_H * 2 * 4;           // fp32 partial output sums [128][2]
constexpr int TC_OFF_BAR = TC_OFF_PART + TC_ROWS * 2 * 4;       // mbarriers: full[5], empty[5], W1, A1, L1, A2 (first half), H0, H1, A2 (second half)
constexpr int TC_OFF_TMEM = TC_OFF_BAR + 18 * 8;
constexpr int TC_SMEM_BYTES = TC_OFF_TMEM + 16;
static_assert(TC_A1_BYTES + TC_W1_BYTES <= TC_A2_BYTES && TC_SMEM_BYTES <= 227 * 1024, "shared memory map");

struct PolicyTC {
  __nv_bfloat16* w1c = nullptr;  // [4 k-cores][64 n-cores][8][8]
  __nv_bfloat16* w2c = nullptr;  // [2 N halves][16 K slabs][4 k-cores][32 n-cores][8][8]
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
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }   // the eight gather / epilogue warps
// non-blocking phase test (mbarrier.try_wait may park the thread; this one never does)
__device__ __forceinline__ uint32_t mbar_test(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok;
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
// One lane of a converged warp.  The warp-wide roles below (producer, MMA issuer) run their loops with all 32 lanes and elect a lane only around
// the tcgen05 / TMA instructions: those take their operands from the uniform datapath, and issued from inside an `if (lane == 0)` region the
// compiler wraps every one of them in a waterfall loop of R2UR moves (measured: the tensor pipe idle half of the time).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
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
__device__ __forceinline__ void tmem_ld32_issue(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
        "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
        "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
// tcgen05.wait::ld, written so that every use of v[] depends on it (the registers are only valid after the wait)
__device__ __forceinline__ void tmem_ld_wait(uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.wait::ld.sync.aligned;"
      : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]), "+r"(v[9]), "+r"(v[10]),
        "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]), "+r"(v[16]), "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]),
        "+r"(v[21]), "+r"(v[22]), "+r"(v[23]), "+r"(v[24]), "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]), "+r"(v[30]), "+r"(v[31])
      :
      : "memory");
}
// 32 registers per thread -> 32 lanes x 32 consecutive fp32 columns
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      :
      : "r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
        "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]),
        "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// ReLU + round to bf16 + pack in one instruction: low half = relu(lo), high half = relu(hi)
__device__ __forceinline__ uint32_t pack_relu_bf16(uint32_t lo, uint32_t hi) {
  uint32_t d;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(__uint_as_float(hi)), "f"(__uint_as_float(lo)));
  return d;
}
// packed fp32 pair arithmetic (sm_100): acc.{lo,hi} += a.{lo,hi} * w.{lo,hi}
__device__ __forceinline__ uint64_t pack_f32x2(float lo, float hi) {
  uint64_t d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi));
  return d;
}
__device__ __forceinline__ void fma_f32x2(uint64_t& acc, uint64_t a, uint64_t w) { asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(a), "l"(w)); }
__device__ __forceinline__ float sum_f32x2(uint64_t v) {
  float lo, hi;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
  return lo + hi;
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);  // .x = lo (low half), .y = hi
  return *reinterpret_cast<uint32_t*>(&t);
}

}  // namespace tc

// Developer timeline (tools/policy_tc_timeline.cu builds with -DTFODE_TC_TIMELINE): clock64 at the role hand-overs of every CTA
#ifdef TFODE_TC_TIMELINE
__device__ long long g_tc_timeline[2048 * 16];
__device__ long long g_tc_slabs[64 * 4];   // CTA 0: per slab {TMA issued, full-wait begins, full-wait ends, MMAs issued}
#define TC_MARK(i) (g_tc_timeline[(blockIdx.x & 2047) * 16 + (i)] = clock64())
#define TC_SLAB_MARK(s, j) ((void)0)
__device__ unsigned g_tc_loop[32 * 8];     // CTA 0: clock() inside the MMA issue loop, per slab
#ifdef TFODE_TC_LOOPMARKS   // five clock reads per slab: they cost about 80 clk of the 330 they measure
#define TC_LOOP_DECL __shared__ unsigned tc_loop_sm[32 * 8];
#define TC_LMARK(s, j) (tc_loop_sm[(s) * 8 + (j)] = (unsigned)clock())
#define TC_LOOP_FLUSH do { if (blockIdx.x == 0) for (int i_ = threadIdx.x & 31; i_ < 32 * 8; i_ += 32) g_tc_loop[i_] = tc_loop_sm[i_]; } while (0)
#else
#define TC_LOOP_DECL
#define TC_LMARK(s, j) ((void)0)
#define TC_LOOP_FLUSH ((void)0)
#endif
#else
#define TC_LOOP_DECL
#define TC_LMARK(s, j) ((void)0)
#define TC_LOOP_FLUSH ((void)0)
#define TC_MARK(i) ((void)0)
#define TC_SLAB_MARK(s, j) ((void)0)
#endif

// in_all: [*,17] scaled observations; list/n_list: parked environments (device) or nullptr for direct rows [0, m_direct).
// b1 is not read here: policy_tc_upload folds it into w1c (rows k = 17, 18); the parameter keeps the launch signature of the fp32 kernel.
__global__ void __launch_bounds__(TC_THREADS, 1)
policy_tc_kernel(const float* __restrict__ in_all, const int* __restrict__ list, const int* __restrict__ n_list, int m_direct,
                 const __nv_bfloat16* __restrict__ w1c, const __nv_bfloat16* __restrict__ w2c, const float* __restrict__ b1,
                 const float* __restrict__ b2, const float* __restrict__ W3, const float* __restrict__ b3, EnvRec* rec, float* act_direct) {
  using namespace tc;
  extern __shared__ __align__(1024) uint8_t smem[];
  const int m = list ? *n_list : m_direct;
  const int row0 = blockIdx.x * TC_ROWS;
  if (row0 >= m) return;                       // uniform per CTA, before any allocation
  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  if (t == 0) TC_MARK(0);

  const uint32_t sA2 = smem_u32(smem + TC_OFF_A2), sRing = smem_u32(smem + TC_OFF_RING), sA1 = smem_u32(smem + TC_OFF_A1), sW1 = smem_u32(smem + TC_OFF_W1);
  const uint32_t barFull = smem_u32(smem + TC_OFF_BAR), barEmpty = barFull + 8 * TC_NSLOT, barW1 = barFull + 80, barA1 = barFull + 88, barL1 = barFull + 96,
                 barA2 = barFull + 104, barH0 = barFull + 112, barH1 = barFull + 120, barA2b = barFull + 128;
  float* sb2 = reinterpret_cast<float*>(smem + TC_OFF_B2);
  float* sW3 = reinterpret_cast<float*>(smem + TC_OFF_W3);
  float* sPart = reinterpret_cast<float*>(smem + TC_OFF_PART);
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + TC_OFF_TMEM);

  if (t == 0) {
    for (int i = 0; i < TC_NSLOT; i++) { mbar_init(barFull + 8 * i, 1); mbar_init(barEmpty + 8 * i, 1); }
    mbar_init(barW1, 1); mbar_init(barA1, TC_ROWS); mbar_init(barL1, 1); mbar_init(barA2, 256); mbar_init(barA2b, 256); mbar_init(barH0, 1); mbar_init(barH1, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 9) {  // one warp owns the TMEM allocation (all 512 columns: 128 lanes x 512 fp32 accumulators)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem + TC_OFF_TMEM)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  if (t == 0) TC_MARK(1);

  if (warp == 8) {
    // ---- producer: W1, then the 32 W2 slabs through the ring
    if (elect_one()) {
      mbar_expect_tx(barW1, TC_W1_BYTES);
      tma_bulk_g2s(sW1, w1c, TC_W1_BYTES, barW1);
    }
    __syncwarp();
    int slot = 0; uint32_t phase = 1;            // parity of the "slot is free" phase the producer waits for; nothing to wait for in round 0
#pragma unroll 1
    for (int s = 0; s < TC_NSLAB; s++) {
      if (s >= TC_NSLOT) mbar_wait(barEmpty + 8 * slot, phase);
      const int gslab = s;                            // slabs 0-15: N half 0, K ascending; 16-31: N half 1 (every CTA in the same order: a row's
                                                      // action does not depend on where in the batch it sits; a per-CTA rotation was tried against
                                                      // L2 hot-spotting and changed nothing)
      if (elect_one()) {
        TC_SLAB_MARK(s, 0);
        mbar_expect_tx(barFull + 8 * slot, TC_SLAB_BYTES);
        tma_bulk_g2s(sRing + slot * TC_SLAB_BYTES, reinterpret_cast<const uint8_t*>(w2c) + (size_t)gslab * TC_SLAB_BYTES, TC_SLAB_BYTES, barFull + 8 * slot);
      }
      __syncwarp();
      if (++slot == TC_NSLOT) { slot = 0; phase ^= 1u; }
    }
    if (lane == 0) TC_MARK(11);
  } else if (warp == 9) {
    // ---- MMA issuer
    mbar_wait(barW1, 0);
    mbar_wait(barA1, 0);
    tc_fence_after();
    if (lane == 0) TC_MARK(2);
    // layer 1: H1 = X(128x32) . W1^T, N in two halves of 256 columns
    {
      const uint64_t ad0 = make_desc(sA1, TC_ROWS * 16, 128), bd0 = make_desc(sW1, TC_H * 16, 128);
      if (elect_one()) {
#pragma unroll
        for (int ks = 0; ks < TC_K1 / 16; ks++)
#pragma unroll
          for (int h = 0; h < 2; h++)
            mma_bf16(tmem + h * 256, ad0 + (uint64_t)((ks * 2 * (TC_ROWS * 16)) >> 4), bd0 + (uint64_t)((ks * 2 * (TC_H * 16) + h * (256 / 8) * 128) >> 4), IDESC_128x256, ks > 0);
        mma_commit(barL1);
      }
      __syncwarp();
    }
    // layer 2: H2 = relu(H1)(128x512) . W2^T; N half 0 over all of K, then N half 1.  It starts when epilogue 1 is through columns 0-255 (the
    // first 8 K slabs of the A operand, and the accumulator columns of N half 0); the other half of epilogue 1 runs under those 16 MMAs.
    mbar_wait(barA2, 0);
    tc_fence_after();
    if (lane == 0) TC_MARK(3);
    const uint64_t ad0 = make_desc(sA2, TC_ROWS * 16, 128), bd0 = make_desc(sRing, TC_SLAB_N * 16, 128);
    // No tcgen05.fence in this loop: the slab's bytes reach the MMA through the mbarrier (TMA complete_tx -> wait -> MMA, all in the async proxy).
    // The whole loop runs in ONE elected lane: per slab it is a barrier poll, two MMA issues and a commit, ~50 clk of issue latency each, and
    // they have to stay under the 2 x 128 clk the tensor pipe needs for the slab.  Electing a lane per slab (and reconverging the warp after it)
    // cost another ~200 clk per slab and left the pipe idle a third of the time (tools/policy_tc_timeline.cu).
    TC_LOOP_DECL
    if (elect_one()) {
      // Fully unrolled: slot, phase and N half are literals, and the barrier test of slab s+1 is issued before the MMAs of slab s so that its
      // shared-memory round trip (56 clk) is off the issue path.
      uint32_t ready = mbar_test(barFull, 0);
#pragma unroll
      for (int s = 0; s < TC_NSLAB; s++) {
        const int slot = s % TC_NSLOT, h = s >> 4, kslab = s & 15;
        const uint32_t phase = (uint32_t)(s / TC_NSLOT) & 1u;
        TC_LMARK(s, 0);
        if (s == 8) { mbar_wait(barA2b, 0); tc_fence_after(); }   // K >= 256 of the A operand and the accumulator columns of N half 1
        if (!ready) mbar_wait(barFull + 8 * slot, phase);   // the slab's bytes reach the MMA through the mbarrier (TMA complete_tx -> wait -> MMA, all in the async proxy)
        if (s + 1 < TC_NSLAB) ready = mbar_test(barFull + 8 * ((s + 1) % TC_NSLOT), (uint32_t)((s + 1) / TC_NSLOT) & 1u);
        TC_LMARK(s, 1);
        const uint64_t ad = ad0 + (uint64_t)(kslab * ((TC_SLAB_K / 8) * (TC_ROWS * 16) >> 4));
        const uint64_t bd = bd0 + (uint64_t)(slot * (TC_SLAB_BYTES >> 4));
#pragma unroll
        for (int ks = 0; ks < TC_SLAB_K / 16; ks++) {  // the accumulators start from b2 (epilogue 1 put it there): always accumulate
          mma_bf16(tmem + h * 256, ad + (uint64_t)(ks * (2 * (TC_ROWS * 16) >> 4)), bd + (uint64_t)(ks * (2 * (TC_SLAB_N * 16) >> 4)), IDESC_128x256, 1u);
          TC_LMARK(s, 3 + ks);
        }
        mma_commit(barEmpty + 8 * slot);       // arrives when the MMAs above have finished reading the slot
        if (s == TC_NSLAB / 2 - 1) { mma_commit(barH0); TC_MARK(12); }
        if (s == TC_NSLAB - 1) { mma_commit(barH1); TC_MARK(4); }
        TC_LMARK(s, 5);
      }
    }
    __syncwarp();
    TC_LOOP_FLUSH;
  } else {
    // ---- gather and epilogues: lane quarter q = TMEM lanes 32 q .., column group g
    const int q = warp & 3, g = warp >> 2;
    const int r = q * 32 + lane;               // row of the tile this thread owns in TMEM
    const int row = row0 + r;
    const bool valid = row < m;
    const int src = valid ? (list ? list[row] : row) : -1;
    if (t < TC_ROWS) {
      // A1: row t, K = 32 (17 observations, zero padded), 4 k-cores of 16 bytes  (t == r for warps 0-3)
      float x[TC_K1];
#pragma unroll
      for (int k = 0; k < TC_K1; k++) x[k] = (valid && k < 17) ? in_all[(size_t)src * 17 + k] : ((k == 17 || k == 18) ? 1.0f : 0.0f);   // k = 17, 18 multiply the two bf16 pieces of b1
#pragma unroll
      for (int kc = 0; kc < TC_K1 / 8; kc++) {
        uint4 qv = make_uint4(pack_bf16(x[kc * 8], x[kc * 8 + 1]), pack_bf16(x[kc * 8 + 2], x[kc * 8 + 3]), pack_bf16(x[kc * 8 + 4], x[kc * 8 + 5]),
                              pack_bf16(x[kc * 8 + 6], x[kc * 8 + 7]));
        *reinterpret_cast<uint4*>(smem + TC_OFF_A1 + kc * (TC_ROWS * 16) + t * 16) = qv;
      }
      fence_proxy_async();     // generic-proxy stores of A1 must be visible to the tensor core (async proxy)
      mbar_arrive(barA1);
    } else {
      // biases and the output layer, fp32
      for (int i = t - TC_ROWS; i < TC_H; i += TC_ROWS) { sb2[i] = b2[i]; sW3[i] = W3[2 * i]; sW3[TC_H + i] = W3[2 * i + 1]; }
    }
    epi_bar_sync();
    const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);

    // ---- epilogue 1: ReLU + bf16 (b1 came in through the MMA), written as the layer-2 A operand; columns 0-255 first (group g takes 128 of
    //      them), hand-over to the MMA warp, then columns 256-511
    mbar_wait(barL1, 0);
    tc_fence_after();
    if (t == 0) TC_MARK(5);
#pragma unroll 1
    for (int ci = 0; ci < 8; ci += 2) {          // two chunks of 32 columns in flight per round trip to TMEM (four at once measured slower)
      const int c = (ci >> 2) * 8 + g * 4 + (ci & 3);
      if (ci == 4) {                              // columns 0-255 are drained, refilled with b2 and in shared memory as bf16: layer 2 may start
        tmem_wait_st();
        fence_proxy_async();
        tc_fence_before();
        mbar_arrive(barA2);
      }
      uint32_t v[2][32], bias[2][32];
      tmem_ld32_issue(lane_base + c * 32, v[0]);
      tmem_ld32_issue(lane_base + c * 32 + 32, v[1]);
#pragma unroll
      for (int u = 0; u < 2; u++)
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const uint4 bb = *reinterpret_cast<const uint4*>(sb2 + (c + u) * 32 + i);
          bias[u][i] = bb.x; bias[u][i + 1] = bb.y; bias[u][i + 2] = bb.z; bias[u][i + 3] = bb.w;
        }
      tmem_ld_wait(v[0]); tmem_ld_wait(v[1]);
#pragma unroll
      for (int u = 0; u < 2; u++) {
#pragma unroll
        for (int gg = 0; gg < 4; gg++) {
          uint4 qv = make_uint4(pack_relu_bf16(v[u][gg * 8], v[u][gg * 8 + 1]), pack_relu_bf16(v[u][gg * 8 + 2], v[u][gg * 8 + 3]),
                                pack_relu_bf16(v[u][gg * 8 + 4], v[u][gg * 8 + 5]), pack_relu_bf16(v[u][gg * 8 + 6], v[u][gg * 8 + 7]));
          *reinterpret_cast<uint4*>(smem + TC_OFF_A2 + ((c + u) * 4 + gg) * (TC_ROWS * 16) + r * 16) = qv;
        }
        tmem_st32(lane_base + (c + u) * 32, bias[u]);   // these 32 columns are read: they become layer-2 accumulators and start from b2
      }
    }
    tmem_wait_st();
    fence_proxy_async();
    tc_fence_before();       // the tcgen05.ld reads of H1 are ordered before the MMAs that overwrite those columns
    mbar_arrive(barA2b);
    if (t == 0) TC_MARK(6);

    // ---- epilogue 2: ReLU (b2 was the accumulators' start value), output layer 512 -> 2 in fp32; N half 0 while half 1 is in the tensor pipe, group g takes 128 columns of each
    uint64_t acc0 = 0ull, acc1 = 0ull;   // (even columns, odd columns) partial sums of output 0 / output 1
#pragma unroll 1
    for (int half = 0; half < 2; half++) {
      mbar_wait(half ? barH1 : barH0, 0);
      tc_fence_after();
      if (t == 0) TC_MARK(half ? 9 : 7);
      const int c0 = half * 8 + g * 4;
      uint32_t v[4][32];
#pragma unroll
      for (int ci = 0; ci < 4; ci++) tmem_ld32_issue(lane_base + (c0 + ci) * 32, v[ci]);
#pragma unroll
      for (int ci = 0; ci < 4; ci++) tmem_ld_wait(v[ci]);
#pragma unroll
      for (int ci = 0; ci < 4; ci++) {
        const int c = c0 + ci;
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const ulonglong2 w0 = *reinterpret_cast<const ulonglong2*>(sW3 + c * 32 + i), w1 = *reinterpret_cast<const ulonglong2*>(sW3 + TC_H + c * 32 + i);
          const uint64_t a01 = pack_f32x2(fmaxf(__uint_as_float(v[ci][i]), 0.0f), fmaxf(__uint_as_float(v[ci][i + 1]), 0.0f));
          const uint64_t a23 = pack_f32x2(fmaxf(__uint_as_float(v[ci][i + 2]), 0.0f), fmaxf(__uint_as_float(v[ci][i + 3]), 0.0f));
          fma_f32x2(acc0, a01, w0.x); fma_f32x2(acc1, a01, w1.x);
          fma_f32x2(acc0, a23, w0.y); fma_f32x2(acc1, a23, w1.y);
        }
      }
    }
    const float s0 = sum_f32x2(acc0), s1 = sum_f32x2(acc1);
    if (t == 0) TC_MARK(8);
    if (g == 1) { sPart[2 * r] = s0; sPart[2 * r + 1] = s1; }
    epi_bar_sync();
    if (g == 0 && valid) {
      const float a0 = tfn::tanh(s0 + sPart[2 * r] + b3[0]), a1 = tfn::tanh(s1 + sPart[2 * r + 1] + b3[1]);
      if (list) { EnvRec& e = rec[src]; e.c.policy_act[0] = a0; e.c.policy_act[1] = a1; e.c.parked = 2; }
      else { act_direct[(size_t)row * 2] = a0; act_direct[(size_t)row * 2 + 1] = a1; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (t == 0) TC_MARK(10);
  if (warp == 9) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
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
// W2 for the ring: [N half h][K slab of 32] -> one canonical operand of R = 256 rows (n) x 32 k, 16 KB contiguous per slab
inline void tc_pack_w2_slabs(const float* W, std::vector<uint16_t>& out) {
  out.assign((size_t)TC_H * TC_H, 0);
  for (int k = 0; k < TC_H; k++)
    for (int n = 0; n < TC_H; n++) {
      const int h = n / TC_SLAB_N, nl = n % TC_SLAB_N, slab = k / TC_SLAB_K, kl = k % TC_SLAB_K;
      const size_t base = (size_t)(h * (TC_H / TC_SLAB_K) + slab) * (TC_SLAB_BYTES / 2);
      out[base + (size_t)(kl / 8) * (TC_SLAB_N / 8) * 64 + (size_t)(nl / 8) * 64 + (nl % 8) * 8 + (kl % 8)] = tc_bf16_rn(W[(size_t)k * TC_H + n]);
    }
}
inline int policy_tc_upload(PolicyTC& tcw, const float* w_host, cudaStream_t stream) {
  const size_t H = TC_H;
  const float* W1 = w_host; const float* W2 = w_host + 17 * H + H;
  std::vector<uint16_t> p1, p2;
  tc_pack_operand(W1, 17, TC_K1, p1);
  {  // b1 rides in the K padding of layer 1 as two bf16 pieces (hi + lo: 2^-17 relative), multiplied by the ones the gather puts at k = 17, 18
    const float* b1 = w_host + 17 * H;
    for (int n = 0; n < TC_H; n++) {
      const uint16_t hi = tc_bf16_rn(b1[n]);
      uint32_t u = (uint32_t)hi << 16; float fhi; memcpy(&fhi, &u, 4);
      const uint16_t lo = tc_bf16_rn(b1[n] - fhi);
      const uint16_t piece[2] = {hi, lo};
      for (int j = 0; j < 2; j++) { const int k = 17 + j; p1[(size_t)(k / 8) * (TC_H / 8) * 64 + (size_t)(n / 8) * 64 + (n % 8) * 8 + (k % 8)] = piece[j]; }
    }
  }
  tc_pack_w2_slabs(W2, p2);
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
  policy_tc_kernel<<<blocks, TC_THREADS, TC_SMEM_BYTES, stream>>>(in, list, n_list, m_direct, tcw.w1c, tcw.w2c, b1, b2, W3, b3, rec, act_direct);
  return cudaGetLastError();
}
}  // namespace tfode
