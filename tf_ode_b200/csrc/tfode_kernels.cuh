// step_kernel and the small per-env utility kernels.  See tfode_step.cuh for the design.
#pragma once
#include "../../include/tfode_lcp.h"
#include "tfode_step.cuh"

namespace tfode {

constexpr int WARPS_PER_CTA = 4;   // reset kernel
// Code-size switch (A/B): the reorder helpers as real functions instead of three inlined copies each (measured slower).
#ifdef TFODE_NOINLINE_REORDER
#define TFODE_REORDER_INLINE __noinline__
#else
#define TFODE_REORDER_INLINE __forceinline__
#endif
#define TFODE_PRAGMA_(x) _Pragma(#x)
#define TFODE_UNROLL(n) TFODE_PRAGMA_(unroll n)
#ifndef TFODE_EXP
#define TFODE_EXP 0   // developer timing experiments of the lock-step body (tools/gpu_exp.sh); 0 = product
#endif
#ifndef TFODE_ENVS_PER_CTA
#define TFODE_ENVS_PER_CTA 8
#endif
constexpr int ENVS_PER_CTA = TFODE_ENVS_PER_CTA;    // step kernel: one warp per environment; 16 environments resident per SM (registers and shared memory alike)

// warp-inclusive scan of small counts
__device__ __forceinline__ int warp_incl_scan(int v, int lane) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { int t = __shfl_up_sync(FULLMASK, v, d); if (lane >= d) v += t; }
  return v;
}

// ------------------------------------------------------------------------------------------------
// (re)construction of an environment: everything simulation() builds (src/rover_training_1.cc:41-125)
// body registers of lane b are returned through the references; ctrl block is written by lane 0.
// ------------------------------------------------------------------------------------------------
__device__ void reset_env(WarpSm& S, const StepParams& P, uint32_t env_global, uint32_t episode, int lane,
                          float* pos, float* q, float* lvel, float* avel, float* facc, float* tacc) {
  const Model& M = g_model;
  const TfodeBatchConfig& cfg = P.cfg;
  if (lane == 0) {
    Ctrl& c = S.c;
    const int keep_explore = c.explore, keep_eps = c.episodes_done, keep_done = c.done_call, keep_term = c.term_reason;
    const float ler = c.last_ep_return, lex = c.last_ep_x, ley = c.last_ep_y; const int les = c.last_ep_steps;
    uint32_t rnd[4]; philox(cfg.seed, env_global, episode, 0, 0, rnd);
    double yaw = (double)cfg.step_yaw_deg + (double)cfg.step_yaw_jitter_deg * (double)u11(rnd[0]);
    float ic_start = (float)((double)cfg.ic_start + (double)cfg.ic_start_jitter * (double)u11(rnd[1]));
    float amp = (float)((double)cfg.crawler_amp + (double)cfg.crawler_amp_jitter * (double)u11(rnd[2]));
    float rate = (float)((double)cfg.crawler_rate + (double)cfg.crawler_rate_jitter * (double)u11(rnd[3]));
    unsigned char* z = reinterpret_cast<unsigned char*>(&c);
    for (int i = 0; i < (int)sizeof(Ctrl); i++) z[i] = 0;
    c.explore = keep_explore; c.episodes_done = keep_eps; c.done_call = keep_done; c.term_reason = keep_term;
    c.last_ep_return = ler; c.last_ep_x = lex; c.last_ep_y = ley; c.last_ep_steps = les;
    c.episode = episode; c.ic_start = ic_start; c.step_yaw_deg = (float)yaw; c.cr_amp = amp; c.cr_rate = rate;
    c.ic_period = cfg.ic_period; c.ic_activated = cfg.ic_initially_active != 0;
    for (int i = 0; i < 4; i++) c.wheel_fmax[i] = 25.4f;  // WHEELS_MAX_TORQUE, src/rover_1.cc:229
    c.cr_phase = 1;
    c.cr_dtorque_dt = (double)(2 * amp / cfg.crawler_dt_torque);
    // step box 0: dRFromEulerAngles(0,0,yaw) -> quaternion (ode/object.cc:116-121); fix joint frames (object.cc:155-179)
    double psi = yaw * 3.14159265358979323846 / 180.0;
    double cp, sp; tfn::d_sincos(psi, &sp, &cp);
    double tr = 2 * cp + 1;  // |yaw| < 120 deg
    double s = sqrt(tr + 1), q0 = 0.5 * s, q3 = (-sp - sp) * (0.5 / s);
    double nrm = 1.0 / sqrt(q0 * q0 + q3 * q3); q0 *= nrm; q3 *= nrm;
    S.aux[0] = (float)q0; S.aux[1] = (float)q3;
    c.fix_qrel[0][0] = (float)q0; c.fix_qrel[0][1] = 0; c.fix_qrel[0][2] = 0; c.fix_qrel[0][3] = (float)(-q3);
    c.fix_qrel[1][0] = 1; c.fix_qrel[1][1] = 0; c.fix_qrel[1][2] = 0; c.fix_qrel[1][3] = 0;
    for (int k = 0; k < 3; k++) { c.fix_offset[0][k] = M.pos0[10][k]; c.fix_offset[1][k] = M.pos0[11][k]; }
    // Rover_1_tf ctor: _last_pos = GetPosition() at the construction pose (identity rotation of the main body)
    for (int k = 0; k < 3; k++) c.last_pos[k] = (double)(M.pos0[0][k] + M.getpos_rel[k]);
  }
  __syncwarp();
  if (lane < NB) {
    for (int k = 0; k < 3; k++) { pos[k] = M.pos0[lane][k]; lvel[k] = 0; avel[k] = 0; facc[k] = 0; tacc[k] = 0; }
    for (int k = 0; k < 4; k++) q[k] = M.quat0[lane][k];
    if (lane == 10) { q[0] = S.aux[0]; q[1] = 0; q[2] = 0; q[3] = S.aux[1]; }
  }
  __syncwarp();
}

// lane b publishes its pose/velocity to the shared body table
__device__ __forceinline__ void publish_body(WarpSm& S, int lane, const float* pos, const float* q, const float* R, const float* lvel, const float* avel) {
  if (lane < NB) {
#pragma unroll
    for (int k = 0; k < 3; k++) { S.bpos[lane][k] = pos[k]; S.bvel[lane][k] = lvel[k]; S.bvel[lane][3 + k] = avel[k]; }
#pragma unroll
    for (int k = 0; k < 4; k++) S.bq[lane][k] = q[k];
#pragma unroll
    for (int k = 0; k < 9; k++) S.bR[lane][k] = R[k];
  }
}

// joint kinematic read-outs of the current pose (servo angle/rate, boggie angle, wheel rates)
__device__ __forceinline__ void joint_readouts(WarpSm& S, int lane) {
  if (lane >= 1 && lane <= 6) {
    if (lane <= 2) S.c.jq_angle[lane] = hinge_angle(S, lane);
    S.c.jq_rate[lane] = hinge_rate(S, lane);
  }
}

// ------------------------------------------------------------------------------------------------
__device__ void append_transition(const StepParams& P, const Ctrl& c, uint32_t env_global, bool terminal, float penalty) {
  int slot = atomicAdd(P.n_trans, 1);
  if (slot >= P.trans_cap) return;
  float* row = P.trans + (size_t)slot * TFODE_TRANSITION_WIDTH;
  row[0] = (float)env_global;
  for (int i = 0; i < 17; i++) row[1 + i] = c.pending[i];
  row[18] = c.pending[17]; row[19] = c.pending[18];
  row[20] = (float)(c.pending_reward - (double)penalty); row[21] = terminal ? 1.0f : 0.0f;
  for (int i = 0; i < 17; i++) row[22 + i] = c.pending[20 + i];
  for (int i = 39; i < TFODE_TRANSITION_WIDTH; i++) row[i] = 0;
}

// Rover_1_tf::GetState (src/rover_1_tf.cc:60-77) from the shared body table; lane-0 code
__device__ void get_state17(const WarpSm& S, float* s) {
  const float* R = S.bR[0];
  float vx0 = R[0], vx1 = R[3], vx2 = R[6];  // body x axis in the world
  float direction = tfn::asin(vx1) * 57.29577951308232f;
  if (vx0 < 0) direction = (vx1 > 0 ? 1.0f : -1.0f) * 180.0f - direction;
  s[0] = direction;
  s[1] = S.c.jq_angle[1] * 57.29577951308232f;
  s[2] = tfn::asin(R[7]) * 57.29577951308232f;   // body y axis, z component
  s[3] = tfn::asin(-vx2) * 57.29577951308232f;
  s[4] = S.c.jq_angle[2] * 57.29577951308232f;
  for (int i = 0; i < 12; i++) s[5 + i] = S.c.ft[i];
}

// ------------------------------------------------------------------------------------------------
// Fused actor (TF_model<float>::infer, tf_cpp_binding.cc:70-113; 17 -> 512 relu -> 512 relu -> 2 tanh of
// scripts/rover_training_1.py:26-35) for ONE environment, run by the environment's own warp at its IC tick.
// Lane l owns hidden units l, l+32, ..., l+480: 16 independent accumulators per lane, each summed in input order
// starting from the bias with one fma per weight -- the order of policy_fp32_kernel and of the oracle, so the
// result is bit-identical to both.  Weights stream from L2 (1.09 MB, shared by every environment of the GPU) with
// fully coalesced 128-byte rows; inputs and hidden activations sit in the environment's row scratch (dead
// between the sweeps and the next sub-step's row build) and are read as shared-memory broadcasts.
// Cost: 8.7 k fma + 8.7 k coalesced loads per lane, once per 0.5 s of simulated time (every 500 sub-steps).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float* actor_scratch(WarpSm& S) { return &S.rowJ[0][0]; }   // [0..17) inputs / outputs, [32..544) h1, [544..1056) h2
static_assert(sizeof(float) * MAXROWS * 12 >= sizeof(float) * (32 + 2 * TFODE_MLP_HIDDEN), "actor scratch aliases the row array");
__device__ __noinline__ void actor_fused(WarpSm& S, const float* __restrict__ w, int lane) {
  constexpr int H = TFODE_MLP_HIDDEN, U = H / 32;
  const float* W1 = w; const float* b1 = W1 + 17 * H; const float* W2 = b1 + H; const float* b2 = W2 + H * H;
  const float* W3 = b2 + H; const float* b3 = W3 + H * 2;
  float* x = actor_scratch(S); float* h1 = x + 32; float* h2 = h1 + H;
  float acc[U];
#pragma unroll
  for (int u = 0; u < U; u++) acc[u] = __ldg(b1 + lane + 32 * u);
  for (int i = 0; i < 17; i++) {
    const float xi = x[i];
#pragma unroll
    for (int u = 0; u < U; u++) acc[u] = tfn::fma(xi, __ldg(W1 + i * H + lane + 32 * u), acc[u]);
  }
#pragma unroll
  for (int u = 0; u < U; u++) h1[lane + 32 * u] = acc[u] > 0 ? acc[u] : 0;
  __syncwarp();
#pragma unroll
  for (int u = 0; u < U; u++) acc[u] = __ldg(b2 + lane + 32 * u);
#pragma unroll 2
  for (int i = 0; i < H; i++) {
    const float hi = h1[i];
#pragma unroll
    for (int u = 0; u < U; u++) acc[u] = tfn::fma(hi, __ldg(W2 + i * H + lane + 32 * u), acc[u]);
  }
#pragma unroll
  for (int u = 0; u < U; u++) h2[lane + 32 * u] = acc[u] > 0 ? acc[u] : 0;
  __syncwarp();
  if (lane < 2) {   // output layer: one lane per action, accumulated in input order from the bias
    float s3 = __ldg(b3 + lane);
    for (int i = 0; i < H; i++) s3 = tfn::fma(h2[i], __ldg(W3 + i * 2 + lane), s3);
    x[lane] = tfn::tanh(s3);
  }
  __syncwarp();
}

// ------------------------------------------------------------------------------------------------
// TFODE_SOLVER_DANTZIG: the dWorldStep equivalent (ode/environment.hh:132).  One lane per constraint island copies the
// island's rows out of shared memory into its slice of the per-environment global scratch, runs the dense Dantzig LCP of
// include/tfode_lcp.h -- the very code the CPU oracle runs, so the float results agree bit for bit -- and leaves the
// constraint acceleration of the island's bodies in btmp, where the integrator expects it.  A correctness path: the pivoting
// method is sequential and works on an m x m matrix, which neither fits the per-environment shared memory nor parallelises
// over a warp the way the row sweeps do; QuickStep stays the fast path.
// ------------------------------------------------------------------------------------------------
__device__ __noinline__ void dantzig_env(WarpSm& S, const StepParams& P, int env, int lane) {
  if (lane >= NISL) return;
  const int L = lane;
  const int m = S.isl_m[L], off = S.isl_off[L];
  if (m <= 0) return;   // labels that lost their rows to a merged island have m = 0
  const int m_total = S.c.n_rows;
  float* f = P.lcp_f + ((size_t)env * MAXROWS + off) * LCP_ROW_F;
  int* ii = P.lcp_i + ((size_t)env * MAXROWS + off) * LCP_ROW_I;
  float* J = f; float* rhs = J + 12 * m; float* cfm = rhs + m; float* lo = cfm + m; float* hi = lo + m; float* lam = hi + m; float* work = lam + m;
  int* jb = ii; int* findex = jb + 2 * m; int* iwork = findex + m;
  int k = 0;
  for (int r = 0; r < m_total; r++) {   // the island's rows in row order: joints in creation order, then contacts in generation order
    const int hdr = __float_as_int(ROWF(S, r, HDR));
    const int b1 = hdr & 15, b2 = (hdr >> 4) & 15, fidx = (hdr >> 8) & 255;
    if (S.isl_lab[body_group(b1)] != L) continue;
    for (int c = 0; c < 12; c++) J[k * 12 + c] = S.rowJ[r][c];
    jb[k * 2] = b1; jb[k * 2 + 1] = b2 < NB ? b2 : -1;
    rhs[k] = ROWF(S, r, RHS); cfm[k] = ROWF(S, r, ADCFM); lo[k] = ROWF(S, r, LO); hi[k] = ROWF(S, r, HI);
    findex[k] = fidx != 255 ? k - (r - fidx) : -1;   // a contact's rows are consecutive: the normal row sits r - fidx places before
    k++;
  }
  float invM[NB], invI[NB * 6], fc[NB * 6];
  for (int b = 0; b < NB; b++) {
    invM[b] = S.binvI[b][6];
    for (int c = 0; c < 6; c++) { invI[b * 6 + c] = S.binvI[b][c]; fc[b * 6 + c] = 0.0f; }
  }
  const int st = tfl::dantzig_island<float>(k, J, jb, rhs, cfm, lo, hi, findex, invM, invI, lam, fc, work, iwork);
  if (st) atomicAdd(P.cstats + 2, 1ull);
  for (int b = 0; b < NB; b++) if (S.isl_lab[body_group(b)] == L) for (int c = 0; c < 6; c++) S.btmp[b][c] = fc[b * 6 + c];
}

// ================================================================================================
//                                         THE STEP KERNEL
// One CTA = ENVS_PER_CTA environments.  Collision, row construction, integration and the rover's control
// code run with one warp per environment (lane = body / geom / joint / row).  The Gauss-Seidel sweeps,
// which are sequential inside a constraint island, run in "body-lane" form on half of the CTA's warps (a
// half-warp per environment, lane = rigid body); the other warps prepare the later reorders (pgs_helper).
// ================================================================================================

// Developer-only phase timers (build with -DTFODE_PHASE_TIMING: tools/phase_timing.py); never compiled into the product library.
#ifdef TFODE_PHASE_TIMING
__device__ unsigned long long g_phase_cycles[16];
#define PT_DECL long long pt_t = clock64(); long long pt_acc[6] = {0, 0, 0, 0, 0, 0};
#define PT_MARK(i) { long long t_ = clock64(); pt_acc[i] += t_ - pt_t; pt_t = t_; }
#define PT_FLUSH if (threadIdx.x == 0) { for (int i_ = 0; i_ < 6; i_++) atomicAdd(&g_phase_cycles[i_], (unsigned long long)pt_acc[i_]); }
#define PC_DECL long long pc_t = clock64();
#define PC_MARK(i) { long long t_ = clock64(); if (threadIdx.x == 0) atomicAdd(&g_phase_cycles[12 + i], (unsigned long long)(t_ - pc_t)); pc_t = t_; }
#define PG_DECL long long pg_t = clock64();
#define PG_MARK(i) { long long t_ = clock64(); if (threadIdx.x == 0) atomicAdd(&g_phase_cycles[8 + i], (unsigned long long)(t_ - pg_t)); pg_t = t_; }
#else
#define PC_DECL
#define PC_MARK(i)
#define PG_DECL
#define PG_MARK(i)
#define PT_DECL
#define PT_MARK(i)
#define PT_FLUSH
#endif

// ------------------------------------------------------------------------------------------------
// Gauss-Seidel sweeps, "body-lane" form.  A half-warp owns one environment and lane b of it owns rigid body b:
// the constraint acceleration fc of the body, its world inverse inertia and inverse mass stay in that lane's
// registers for the whole solve.  A row touches at most two bodies, so its relaxation is carried out by the
// two lanes of its bodies: each forms its half of J.fc, one shuffle completes the 12-wide dot product, both
// compute the same delta and each updates its own fc.
//
// Order.  ODE relaxes the rows of an island strictly one after the other, in the (randomly permuted) order[]
// of that island.  Two rows that share no body commute exactly -- neither reads anything the other writes --
// so the sequential result is reproduced bit for bit when every BODY sees its own rows in the island's order:
// each lane walks the queue of the rows that touch its body (the sub-sequence of order[] it takes part in)
// and a row fires as soon as it is at the head of the queues of both of its bodies.  The number of
// dependent steps per sweep drops from the island's row count to the longest per-body chain, and the
// islands of an environment (and the two environments of a warp) advance concurrently.
// ------------------------------------------------------------------------------------------------
// Queue entry (16 bits, decoded once per reorder instead of once per lock-step):
//   row | is1 << 7 | partner body << 8 (the lane's own body for a one-body row) | solo << 12 | fo << 13
// fo = distance back to the normal row of a friction row (contact rows are laid out normal, friction 1, friction 2), 0 otherwise.
struct BodyRow { int meta; float J[6]; float a[3]; float rhs, Ad, Adcfm, lo, hi; };
constexpr int BR_IS1 = 1 << 7, BR_SOLO = 1 << 12;
static_assert(MAXROWS <= 128, "queue entries hold the row index in 7 bits");
__device__ __forceinline__ int pgs_make_entry(int row, int hdr, int b) {
  const int b1 = hdr & 15, b2 = (hdr >> 4) & 15, fidx = (hdr >> 8) & 255;
  const bool is1 = b1 == b, solo = b2 >= NB;
  const int pb = solo ? b : (is1 ? b2 : b1), fo = fidx != 255 ? row - fidx : 0;
  return row | (is1 ? BR_IS1 : 0) | (pb << 8) | (solo ? BR_SOLO : 0) | (fo << 13);
}
// appends the entry of a row (packed pair pk, see RS_ENT) to the queue of body b if the row touches b
__device__ __forceinline__ void pgs_push(unsigned short*& qw, int pk, int b) {
  const bool is1 = ((pk >> 24) & 15) == b, is2 = ((pk >> 8) & 15) == b;
  const int ent = is1 ? (pk & 0xffff) : ((pk >> 16) & 0xffff);
  if (is1 || is2) *qw = (unsigned short)ent;
  qw += (is1 || is2) ? 1 : 0;
}
// static part of a row for this lane's body; lam = the row's lambda as stored right now
__device__ __forceinline__ void pgs_load_row(const WarpSm& S, int entry, const float* iI, BodyRow& r, float& lam) {
  const int row = entry & 127;
  const float2* Jp = reinterpret_cast<const float2*>(&S.rowJ[row][(entry & BR_IS1) ? 0 : 6]);
  const float2 j0 = Jp[0], j1 = Jp[1], j2 = Jp[2];
#if TFODE_EXP == 11   // sensitivity experiment (results unchanged): the three Jacobian loads once more
  { const unsigned ja = (unsigned)__cvta_generic_to_shared(Jp); float d0, d1, d2, d3, d4, d5;
    asm volatile("ld.shared.v2.f32 {%0,%1}, [%6]; ld.shared.v2.f32 {%2,%3}, [%6+8]; ld.shared.v2.f32 {%4,%5}, [%6+16];" : "=f"(d0), "=f"(d1), "=f"(d2), "=f"(d3), "=f"(d4), "=f"(d5) : "r"(ja));
    asm volatile("" ::"f"(d0), "f"(d2), "f"(d4)); }
#endif
#ifdef TFODE_ROWS_AOS
  const float4 s0 = *reinterpret_cast<const float4*>(&S.rowS[row][0]);       // rhs Ad Adcfm lo
  const float2 s1 = *reinterpret_cast<const float2*>(&S.rowS[row][RS_HI]);   // hi lambda
  static_assert(RS_LAM == RS_HI + 1 && (RS_HI % 2) == 0, "hi and lambda travel in one 8-byte load");
#else
  const float4 s0 = S.rowC[row];    // rhs Ad Adcfm lo
  const float2 s1 = S.rowHL[row];   // hi lambda
#if TFODE_EXP == 16   // sensitivity experiment: the two row-scalar loads once more
  { const unsigned ca = (unsigned)__cvta_generic_to_shared(&S.rowC[row]), ha = (unsigned)__cvta_generic_to_shared(&S.rowHL[row]); float d0, d1, d2, d3, d4, d5;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%6]; ld.shared.v2.f32 {%4,%5}, [%7];" : "=f"(d0), "=f"(d1), "=f"(d2), "=f"(d3), "=f"(d4), "=f"(d5) : "r"(ca), "r"(ha));
    asm volatile("" ::"f"(d0), "f"(d4)); }
#endif
#endif
  r.meta = entry;
  r.hi = s1.x; lam = s1.y;
  r.J[0] = j0.x; r.J[1] = j0.y; r.J[2] = j1.x; r.J[3] = j1.y; r.J[4] = j2.x; r.J[5] = j2.y;
  r.rhs = s0.x; r.Ad = s0.y; r.Adcfm = s0.z; r.lo = s0.w;
  // angular part of M^-1 J^T for this body (constant during the sweeps)
  r.a[0] = tfn::dot3(iI[0], iI[1], iI[2], r.J[3], r.J[4], r.J[5]);
  r.a[1] = tfn::dot3(iI[1], iI[3], iI[4], r.J[3], r.J[4], r.J[5]);
  r.a[2] = tfn::dot3(iI[2], iI[4], iI[5], r.J[3], r.J[4], r.J[5]);
}

__device__ __forceinline__ int first_body_of_group(int g) { return g == 0 ? 0 : (g == 1 ? 4 : (g == 2 ? 5 : (g == 3 ? 10 : 11))); }

constexpr int PGS_WARPS = ENVS_PER_CTA / 2;
static_assert(PGS_WARPS <= 15, "one named barrier per warp pair");

// Named barrier over 64 threads: sweep warp p and its helper warp PGS_WARPS + p (barrier 0 stays with __syncthreads).
__device__ __forceinline__ void pair_barrier(int p) { asm volatile("bar.sync %0, 64;" ::"r"(1 + p) : "memory"); }

// The island's order array after reorder r.  ODE shuffles the same array again at every 8th sweep; here the permutations of
// reorders 0, 1 and 2 live in three buffers so that they can be prepared ahead of the sweeps that use them: r = 0 in
// S.order (shuffled by the environment's own warp before the CTA barrier), r = 1 and r >= 2 in bytes 0 and 1 of the rows'
// header slots (RS_HDR is dead once the order is built), written by the pair's helper warp while the sweeps of block 0 run.
__device__ __forceinline__ unsigned char* ord_buffer(WarpSm& S, int off, int r, int& stride) {
  if (r == 0) { stride = 1; return S.order + off; }
  stride = ROW_HDR_STRIDE;
  return reinterpret_cast<unsigned char*>(&ROWF(S, off, HDR)) + (r == 1 ? 0 : 1);
}
// RANDOMLY_REORDER_CONSTRAINTS: Fisher-Yates with ODE's dRandInt on this island's slice of the environment's stream (one lane).
// The draw of the next swap (LCG step, bit folds, reciprocal-multiply modulo) overlaps the shared-memory round trip of this one.
__device__ TFODE_REORDER_INLINE void fy_shuffle(unsigned char* o, int stride, int m, uint32_t seed0, int base, int r) {
  uint32_t sd = lcg_jump(seed0, (uint32_t)(base + r * (m - 1)));
  int sw = lcg_randint(sd, 2u);
  for (int q = 1; q < m; q++) {
    const int sw_next = lcg_randint_magic(sd, (uint32_t)(q + 2), c_fy_magic.v[q + 2]);
    const unsigned char x = o[q * stride], y = o[sw * stride];
    o[q * stride] = y; o[sw * stride] = x;
    sw = sw_next;
  }
}
// Scratch of the sweep phase.  The body tables (bR, bvel: published poses and velocities) are dead between the row build and
// the re-publication after the integration, so the second queue buffer and the schedule hand-over live there.
//   queue buffer 0: the contact list (dead once the rows are built)      -> reorders 0, 2, 3, ...
//   queue buffer 1: bR                                                   -> reorder 1
//   hand-over     : bvel as unsigned[2][17]: fire mask of body lane b, [16] = lock-steps per sweep, for reorders 1 and 2
__device__ __forceinline__ unsigned short* pgs_queue_buffer(WarpSm& S, int which) {
  static_assert(2 * MAXROWS * sizeof(unsigned short) <= sizeof(float) * MAXC * 8, "queues alias the contact list");
  static_assert(2 * MAXROWS * sizeof(unsigned short) <= sizeof(float) * NB * 12, "second queue buffer aliases bR");
  static_assert(2 * 17 * sizeof(unsigned) <= sizeof(float) * NB * 8, "schedule hand-over aliases bvel");
  return which ? reinterpret_cast<unsigned short*>(&S.bR[0][0]) : reinterpret_cast<unsigned short*>(&S.cont[0][0]);
}
__device__ __forceinline__ unsigned* pgs_handover(WarpSm& S, int r) { return reinterpret_cast<unsigned*>(&S.bvel[0][0]) + (r - 1) * 17; }

// One reorder's queue fill + static firing schedule for body lane b (all 32 lanes of the warp take part in the shuffles).
//  * fill: every row carries its two pre-decoded entries (RS_ENT); an entry's partner field names the OTHER body, so this
//    lane's body is body 1 of the row iff the second entry's partner is b, body 2 iff the first entry's partner is b (a
//    one-body row names its own body in both, and its lane takes the first).  Branch-free, four rows per round.
//  * schedule: which lock-step a row fires at depends only on the queues (a row fires when it heads the queues of both of its
//    bodies), not on any value, so the integer part of the lock-step loop is run once per reorder: bit t of `mask0` = this
//    lane relaxes its next queue entry at lock-step t; T = lock-steps per sweep (warp-uniform).
__device__ TFODE_REORDER_INLINE void pgs_fill_sched(const WarpSm& S, const unsigned char* ord, int ostride, int m, int b, int hbase,
                                               unsigned short* myq, int qn, int dummy_ent, unsigned long long& mask0, int& T) {
  unsigned short* qw = myq;
  int k = 0;
  for (; k + 4 <= m; k += 4) {
    const int r0 = ord[k * ostride], r1 = ord[(k + 1) * ostride], r2 = ord[(k + 2) * ostride], r3 = ord[(k + 3) * ostride];
    const int p0 = __float_as_int(ROWF(S, r0, ENT)), p1 = __float_as_int(ROWF(S, r1, ENT));
    const int p2 = __float_as_int(ROWF(S, r2, ENT)), p3 = __float_as_int(ROWF(S, r3, ENT));
    pgs_push(qw, p0, b); pgs_push(qw, p1, b); pgs_push(qw, p2, b); pgs_push(qw, p3, b);
  }
  for (; k < m; k++) pgs_push(qw, __float_as_int(ROWF(S, ord[k * ostride], ENT)), b);
  __syncwarp();
  mask0 = 0ull; T = 0;
  int cur = 0, ent = qn > 0 ? (int)myq[0] : dummy_ent;
  while (__any_sync(FULLMASK, cur < qn)) {
    const bool have = cur < qn;
    const int myrow = have ? (ent & 127) : -1;
    const int prow = __shfl_sync(FULLMASK, myrow, hbase | ((ent >> 8) & 15));
    if (have && ((ent & BR_SOLO) || prow == myrow)) {
      if (T < 64) mask0 |= 1ull << T;
      cur++;
      if (cur < qn) ent = myq[cur];
    }
    T++;
  }
}
// rows of body b in its island (does not depend on the permutation) and the start of its queue inside a queue buffer
__device__ TFODE_REORDER_INLINE void pgs_queue_extent(const WarpSm& S, const unsigned char* ordA, int m, int b, int& qn, int& qoff) {
  qn = 0;
  for (int k = 0; k < m; k++) {
    const int pk = __float_as_int(ROWF(S, ordA[k], ENT));
    if (((pk >> 24) & 15) == b || ((pk >> 8) & 15) == b) qn++;
  }
  int incl = qn;
#pragma unroll
  for (int d = 1; d < 16; d <<= 1) { const int t = __shfl_up_sync(FULLMASK, incl, d, 16); if (b >= d) incl += t; }
  qoff = incl - qn;
}

// Which environment slot of the CTA half `lane >> 4` of pair `pair` works on.  Default: slots 2*pair, 2*pair + 1.
// -DTFODE_SORT_PAIRS: the CTA's environments are ranked by their row count of this sub-step and pair p takes ranks 2p and
// 2p + 1, so that a sweep warp's two environments need about the same number of lock-steps per sweep (the warp runs
// max(T_a, T_b) of them: pairing like with like minimises the lock-steps the CTA issues in total; the longest environment
// sets the phase's length either way).  Sweep warp and helper warp of a pair evaluate the same ranking.
__device__ __forceinline__ int pgs_env_of(const WarpSm* sm, int pair, int lane) {
#ifdef TFODE_SORT_PAIRS
  int w = 0;
  if (lane < ENVS_PER_CTA) w = sm[lane].pgs_on ? sm[lane].c.n_rows : 0;
  int rank = 0;
#pragma unroll
  for (int j = 0; j < ENVS_PER_CTA; j++) { const int wj = __shfl_sync(FULLMASK, w, j); rank += (wj > w || (wj == w && j < lane)) ? 1 : 0; }
  const int want = 2 * pair + (lane >> 4);
  int e = 0;
#pragma unroll
  for (int j = 0; j < ENVS_PER_CTA; j++) { const int rj = __shfl_sync(FULLMASK, rank, j); if (rj == want) e = j; }
  return e;
#else
  return 2 * pair + (lane >> 4);
#endif
}

// Helper warp of pair p: everything reorders 1 and 2 need that does not depend on the solve -- permutation, queues, firing
// schedule -- for the pair's two environments, handed to the sweep warp at two meetings of the pair's named barrier.
__device__ void pgs_helper(WarpSm* sm, int pgs_iters, int p, int lane) {
  const int e = pgs_env_of(sm, p, lane), b = lane & 15, hbase = lane & 16;
  WarpSm& S = sm[e];
  const bool on = S.pgs_on != 0 && b < NB;
  int off = 0, m = 0, base = 0; bool leader = false;
  if (on) {
    const int g = body_group(b), L = S.isl_lab[g];
    off = S.isl_off[L]; m = S.isl_m[L]; base = S.isl_base[L];
    leader = (b == first_body_of_group(g)) && (L == g);
  }
  int sa, sb, sc;
  unsigned char* A = ord_buffer(S, off, 0, sa); unsigned char* B = ord_buffer(S, off, 1, sb); unsigned char* C = ord_buffer(S, off, 2, sc);
  if (leader && m > 0) {
    for (int k = 0; k < m; k++) B[k * sb] = A[k * sa];
    if (m > 1) fy_shuffle(B, sb, m, S.c.lcg_seed, base, 1);
    if (pgs_iters > 16) {
      for (int k = 0; k < m; k++) C[k * sc] = B[k * sb];
      if (m > 1) fy_shuffle(C, sc, m, S.c.lcg_seed, base, 2);
    }
  }
  int qn, qoff;
  pgs_queue_extent(S, A, m, b, qn, qoff);
  __syncwarp();
  const int dummy_ent = (b << 8) | BR_SOLO;
  unsigned long long mask0; int T;
  pgs_fill_sched(S, B, sb, m, b, hbase, pgs_queue_buffer(S, 1) + qoff, qn, dummy_ent, mask0, T);
  if (on) { unsigned* h = pgs_handover(S, 1); h[b] = (unsigned)mask0; if (b == 0) h[16] = (unsigned)T; }
  __syncwarp();
  pair_barrier(p);   // reorder 1 is ready; the sweep warp has left queue buffer 0
  if (pgs_iters > 16) {
    pgs_fill_sched(S, C, sc, m, b, hbase, pgs_queue_buffer(S, 0) + qoff, qn, dummy_ent, mask0, T);
    if (on) { unsigned* h = pgs_handover(S, 2); h[b] = (unsigned)mask0; if (b == 0) h[16] = (unsigned)T; }
    __syncwarp();
    pair_barrier(p);   // reorder 2 is ready
  }
}

template <int LSU>   // unroll factor of the lock-step loop
__device__ void pgs_bodylane(WarpSm* sm, int pgs_iters, int wid, int lane) {
  const int e = pgs_env_of(sm, wid, lane), b = lane & 15;
  int hbase = lane & 16;
  asm volatile("" : "+r"(hbase));   // (same reason as s_addr below)
  // The environment's shared-memory address is made opaque: under the 128-register cap ptxas otherwise rematerialises it
  // from %tid / %cluster_ctaid (two S2R, ~40 cycles each) inside the lock-step loop instead of keeping it in a register.
  uint32_t s_addr = (uint32_t)__cvta_generic_to_shared(&sm[e]);
  asm volatile("" : "+r"(s_addr));
  WarpSm& S = *reinterpret_cast<WarpSm*>(__cvta_shared_to_generic(s_addr));
  const bool on = S.pgs_on != 0 && b < NB;
  int off = 0, m = 0, base = 0; bool leader = false;
  float fc[6] = {0, 0, 0, 0, 0, 0}, iI[6] = {0, 0, 0, 0, 0, 0}, invM = 0;
  uint32_t seed0 = 0;
  if (on) {
    const int g = body_group(b), L = S.isl_lab[g];
    off = S.isl_off[L]; m = S.isl_m[L]; base = S.isl_base[L];
    leader = (b == first_body_of_group(g)) && (L == g);
    const float4 ia = *reinterpret_cast<const float4*>(&S.binvI[b][0]); const float4 ib = *reinterpret_cast<const float4*>(&S.binvI[b][4]);
    iI[0] = ia.x; iI[1] = ia.y; iI[2] = ia.z; iI[3] = ia.w; iI[4] = ib.x; iI[5] = ib.y; invM = ib.z;
    seed0 = S.c.lcg_seed;
  }
  int ostride = 1;
  unsigned char* ord = S.order + off;   // reorder 0: shuffled by the environment's own warp before the CTA barrier
  PG_DECL
#ifdef TFODE_PHASE_TIMING
  int nsteps = 0;
#endif
  int qn, qoff;
  pgs_queue_extent(S, ord, m, b, qn, qoff);
  unsigned short* myq = pgs_queue_buffer(S, 0) + qoff;

  PG_MARK(0)
  const int dummy_ent = (b << 8) | BR_SOLO;   // idle slot: row 0, one-body, partner = self, never fires
  for (int it0 = 0; it0 < pgs_iters; it0 += 8) {
    const int nsw = min(8, pgs_iters - it0);   // sweeps until the next reorder
    unsigned long long mask0 = 0ull; int T = 0;
    {  // RANDOMLY_REORDER_CONSTRAINTS: the permutations of reorders 0..2 are ready in their buffers (ord_buffer); later ones
       // (pgs_iters > 24) are shuffled here, in place on the last buffer
      PG_MARK(3)
      const int rr = it0 >> 3;
      if (rr == 1 || rr == 2) {
        pair_barrier(wid);   // the helper warp finished this reorder long ago (permutation, queues, schedule)
        myq = pgs_queue_buffer(S, rr == 1 ? 1 : 0) + qoff;
        const unsigned* h = pgs_handover(S, rr);
        mask0 = on ? (unsigned long long)h[b] : 0ull;
        T = __reduce_max_sync(FULLMASK, on ? (int)h[16] : 0);
      } else {
        if (rr >= 3) {
          ord = ord_buffer(S, off, rr, ostride);
          if (leader && m > 1) fy_shuffle(ord, ostride, m, seed0, base, rr);
          __syncwarp();
          myq = pgs_queue_buffer(S, 0) + qoff;
        }
        PG_MARK(1)
        pgs_fill_sched(S, ord, ostride, m, b, hbase, myq, qn, dummy_ent, mask0, T);
      }
    }
    PG_MARK(2)
#ifdef TFODE_PHASE_TIMING
    if (threadIdx.x == 0) atomicAdd(&g_phase_cycles[7], (unsigned long long)nsw);
    nsteps += nsw * T;
#endif
    if (T >= 2 && T <= 32) {
      // Lock-steps, software-pipelined over the static schedule: every lane takes every lock-step (idle slots carry a
      // dummy entry and are predicated off), so the body is one basic block -- no readiness test, no vote, no divergent
      // branch -- and the record of lock-step t+1 (queue entry, J, row constants, I^-1 J, lambda) is requested at the
      // top of lock-step t, off the dependent chain fc -> J.fc -> shuffle -> delta -> fc.  lambda of a row is read and
      // written by ONE lane only (the lane of the row's first body, which is also the first body of the friction rows
      // of the same contact) and reaches the partner inside the shuffle round; the one read-after-write hazard the
      // prefetch opens (a friction row right after its own normal row on the same lane) is closed by forwarding.
      const unsigned mask32 = (unsigned)mask0;
      BodyRow rb; float lam, lamf;
      {
        const int e0 = (mask32 & 1u) ? (int)myq[0] : dummy_ent;
        pgs_load_row(S, e0, iI, rb, lam);
        lamf = ROWF(S, (e0 & 127) - ((e0 >> 13) & 3), LAM);
      }
      // one lock-step: `fire` = this lane relaxes rb now; `en` = its queue entry for the next lock-step
      auto lock_step = [&](const bool fire, const int en) {
        BodyRow rn; float lam_n, lamf_n;
        pgs_load_row(S, en, iI, rn, lam_n);
        const int nrow_n = (en & 127) - ((en >> 13) & 3);
        lamf_n = ROWF(S, nrow_n, LAM);
        const int row = rb.meta & 127, pl = hbase | ((rb.meta >> 8) & 15);
        const bool is1 = (rb.meta & BR_IS1) != 0, solo = (rb.meta & BR_SOLO) != 0, fric = (rb.meta >> 13) != 0;
        const float part = tfn::dot6(rb.J, fc);
        const float other = __shfl_sync(FULLMASK, part, pl);
        const float lam_o = __shfl_sync(FULLMASK, lam, pl);
        const float lamf_o = __shfl_sync(FULLMASK, lamf, pl);
#if TFODE_EXP == 13   // sensitivity experiment: two more shuffles (off the dependent chain)
        { const float x1 = __shfl_sync(FULLMASK, rb.rhs, pl), x2 = __shfl_sync(FULLMASK, rb.Ad, pl); asm volatile("" ::"f"(x1), "f"(x2)); }
#endif
        const float lam_c = is1 ? lam : lam_o, lamf_c = is1 ? lamf : lamf_o;
        float lo = rb.lo, hi = rb.hi;
        if (fric) { hi = fabsf(hi * lamf_c); lo = -hi; }
        const float sum = solo ? part : (is1 ? part + other : other + part);
        float delta = tfn::fma(rb.Ad, rb.rhs - sum, -(lam_c * rb.Adcfm));
        float nl = lam_c + delta;
        if (nl < lo) { delta = lo - lam_c; nl = lo; }
        else if (nl > hi) { delta = hi - lam_c; nl = hi; }
#if TFODE_EXP == 15   // sensitivity experiment: four dependent adds of -0 ON the chain fc -> delta -> fc (+16 cycles of latency, same value)
        { float mz = __int_as_float(0x80000000 | (threadIdx.x >> 10)); asm volatile("add.f32 %0, %0, %1; add.f32 %0, %0, %1; add.f32 %0, %0, %1; add.f32 %0, %0, %1;" : "+f"(delta) : "f"(mz)); }
#endif
#if TFODE_EXP == 14   // sensitivity experiment: sixteen ALU instructions off the chain (issue slots only)
        { float z = rb.rhs; asm volatile("add.f32 %0, %0, %0; add.f32 %0, %0, %0; add.f32 %0, %0, %0; add.f32 %0, %0, %0; add.f32 %0, %0, %0; add.f32 %0, %0, %0; add.f32 %0, %0, %0; add.f32 %0, %0, %0;"
                                         "add.f32 %0, %0, %0; add.f32 %0, %0, %0; add.f32 %0, %0, %0; add.f32 %0, %0, %0; add.f32 %0, %0, %0; add.f32 %0, %0, %0; add.f32 %0, %0, %0; add.f32 %0, %0, %0;" : "+f"(z)); asm volatile("" ::"f"(z)); }
#endif
        if (fire && is1) ROWF(S, row, LAM) = nl;
#if TFODE_EXP == 12   // sensitivity experiment: the lambda store once more (same value, same address)
        if (fire && is1) { volatile float* lv = &ROWF(S, row, LAM); *lv = nl; }
#endif
        const float dm = delta * invM;
        const float f0 = tfn::fma(dm, rb.J[0], fc[0]), f1 = tfn::fma(dm, rb.J[1], fc[1]), f2 = tfn::fma(dm, rb.J[2], fc[2]);
        const float f3 = tfn::fma(delta, rb.a[0], fc[3]), f4 = tfn::fma(delta, rb.a[1], fc[4]), f5 = tfn::fma(delta, rb.a[2], fc[5]);
        fc[0] = fire ? f0 : fc[0]; fc[1] = fire ? f1 : fc[1]; fc[2] = fire ? f2 : fc[2];
        fc[3] = fire ? f3 : fc[3]; fc[4] = fire ? f4 : fc[4]; fc[5] = fire ? f5 : fc[5];
        if (fire && nrow_n == row) lamf_n = nl;   // the normal row of the next (friction) row was relaxed just now
        rb = rn; lam = lam_n; lamf = lamf_n;
      };
      const unsigned short* qp0 = myq;
      const int e_first = (mask32 & 1u) ? (int)myq[0] : dummy_ent;
      // The sweeps of this reorder block.  The unroll factor of the lock-step loop trades loop overhead against the loop's instruction
      // footprint (main body x U + remainder body + peeled last step, ~100 instructions each): measured on every workload
      // (profiles/r02_ab_table.md), 2 wins by 6-10 % where sweeps are long (step boxes, wheels in the groove), 4 by 2 % where they are short
      // (plane).  Both loops inside one kernel lose to either (the footprint doubles), so LSU is a template parameter of step_kernel and the
      // launcher picks the instantiation by scene.
#define TFODE_SWEEP_LOOPS(U)                                                                                                  \
      for (int sw_i = 0; sw_i < nsw; sw_i++) {                                                                                \
        unsigned mrem = mask32;                                                                                               \
        const unsigned short* qp = qp0;                                                                                       \
        TFODE_UNROLL(U)                                                                                                       \
        for (int t = 0; t < T - 1; t++) {                                                                                     \
          const bool fire = (mrem & 1u) != 0u;                                                                                \
          qp += fire ? 1 : 0;                                                                                                 \
          mrem >>= 1;                                                                                                         \
          int en = dummy_ent;                                                                                                 \
          if (mrem & 1u) en = *qp;                                                                                            \
          lock_step(fire, en);                                                                                                \
        }                                                                                                                     \
        lock_step((mrem & 1u) != 0u, e_first);   /* last lock-step of the sweep: the next one is lock-step 0 again */         \
      }
      TFODE_SWEEP_LOOPS(LSU)
#undef TFODE_SWEEP_LOOPS
    } else {
    // Fallback (more than 32 lock-steps per sweep, or single-row chains): the same lock-steps with the readiness test done on
    // the fly -- queue entries pre-decoded, lambda owned by one lane and exchanged in the shuffle round, no __syncwarp.
    for (int sw_i = 0; sw_i < nsw; sw_i++) {
    int cur = 0;
    BodyRow rb;
    rb.meta = dummy_ent; rb.rhs = 0; rb.Ad = 0; rb.Adcfm = 0; rb.lo = 0; rb.hi = 0;
#pragma unroll
    for (int k = 0; k < 6; k++) rb.J[k] = 0;
    rb.a[0] = 0; rb.a[1] = 0; rb.a[2] = 0;
    float lam = 0.0f;
    if (qn > 0) pgs_load_row(S, myq[0], iI, rb, lam);
    float lamf = ROWF(S, (rb.meta & 127) - ((rb.meta >> 13) & 3), LAM);
    int ent_next = (qn > 1) ? myq[1] : rb.meta;
    while (__any_sync(FULLMASK, cur < qn)) {
      const bool have = cur < qn;
      const int row = rb.meta & 127, pl = hbase | ((rb.meta >> 8) & 15);
      const bool is1 = (rb.meta & BR_IS1) != 0, solo = (rb.meta & BR_SOLO) != 0, fric = (rb.meta >> 13) != 0;
      const int myrow = have ? row : -1;
      const float part = tfn::dot6(rb.J, fc);
      const int prow = __shfl_sync(FULLMASK, myrow, pl);
      const float other = __shfl_sync(FULLMASK, part, pl);
      const float lam_o = __shfl_sync(FULLMASK, lam, pl);
      const float lamf_o = __shfl_sync(FULLMASK, lamf, pl);
      const float lam_c = is1 ? lam : lam_o, lamf_c = is1 ? lamf : lamf_o;
      const bool ready = have && (solo || prow == myrow);
      float lo = rb.lo, hi = rb.hi;
      if (fric) { hi = fabsf(hi * lamf_c); lo = -hi; }
      const float sum = solo ? part : (is1 ? part + other : other + part);
      float delta = tfn::fma(rb.Ad, rb.rhs - sum, -(lam_c * rb.Adcfm));
      float nl = lam_c + delta;
      if (nl < lo) { delta = lo - lam_c; nl = lo; }
      else if (nl > hi) { delta = hi - lam_c; nl = hi; }
      if (ready) {
        if (is1) ROWF(S, row, LAM) = nl;
        const float dm = delta * invM;
        fc[0] = tfn::fma(dm, rb.J[0], fc[0]); fc[1] = tfn::fma(dm, rb.J[1], fc[1]); fc[2] = tfn::fma(dm, rb.J[2], fc[2]);
        fc[3] = tfn::fma(delta, rb.a[0], fc[3]); fc[4] = tfn::fma(delta, rb.a[1], fc[4]); fc[5] = tfn::fma(delta, rb.a[2], fc[5]);
        cur++;
        if (cur < qn) {
          pgs_load_row(S, ent_next, iI, rb, lam);
          lamf = ROWF(S, (ent_next & 127) - ((ent_next >> 13) & 3), LAM);   // after the store above: sees a normal row relaxed just now
          if (cur + 1 < qn) ent_next = myq[cur + 1];
        }
      }
    }
    }
    }
    __syncwarp();
  }
#ifdef TFODE_PHASE_TIMING
  if (threadIdx.x == 0) atomicAdd(&g_phase_cycles[6], (unsigned long long)nsteps);
#endif
  PG_MARK(3)
  if (on) {
    float* o = S.btmp[b];
    *reinterpret_cast<float4*>(o) = make_float4(fc[0], fc[1], fc[2], fc[3]);
    *reinterpret_cast<float2*>(o + 4) = make_float2(fc[4], fc[5]);
  }
}

#ifndef TFODE_MODEL_SRC
#define TFODE_MODEL_SRC (*P.model)
#endif
// SOLVER: TFODE_SOLVER_QUICKSTEP (the fast path; no trace of the dense solver in its code) or TFODE_SOLVER_DANTZIG.
// LSU: unroll factor of the sweeps' lock-step loop (2 or 4, see pgs_bodylane; chosen per scene by the launcher).
template <int SOLVER, int LSU>
__global__ void __launch_bounds__(ENVS_PER_CTA * 32, 16 / ENVS_PER_CTA) step_kernel(const StepParams P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  WarpSm* const sm = reinterpret_cast<WarpSm*>(smem_raw);
  WarpSm& S = sm[wid];
  const Model& M = TFODE_MODEL_SRC;   // A/B switch: constant memory (g_model) or the global-memory copy (*P.model)
  const TfodeBatchConfig& cfg = P.cfg;

  // Work items.  Plain launch (P.slice_len == 0): CTA b = environment group b for the whole call.  Persistent launch: the grid is one
  // wave of co-resident CTAs pulling (group, slice) items -- slice_len sub-steps of one group of 8 environments -- from an atomic
  // counter in slice-major order, so that a batch of, say, 1.73 waves of groups costs 1.73 waves of time instead of 2.  A group's
  // slices are sequential: item (g, s) waits for slice_done[g] == s.  The item it waits for has a smaller index, was therefore
  // claimed earlier by a CTA that is resident and itself only waits on smaller indices, so the wait always ends.
  for (;;) {
  int group = blockIdx.x, slice = 0;
  if (P.slice_len > 0) {
    __syncthreads();   // every warp is done with the previous item's shared memory
    if (threadIdx.x == 0) {
      const int item = atomicAdd(P.sched, 1);
      int g_ = -1, s_ = 0;
      if (item < P.n_items) {
        g_ = item % P.n_groups; s_ = item / P.n_groups;
        while (ld_acquire(P.slice_done + g_) < s_) __nanosleep(256);
      }
      sm[0].aux[8] = __int_as_float(g_); sm[0].aux[9] = __int_as_float(s_);
    }
    __syncthreads();
    group = __float_as_int(sm[0].aux[8]); slice = __float_as_int(sm[0].aux[9]);
    if (group < 0) return;   // CTA-uniform: the queue is empty
  }
  const int env = group * ENVS_PER_CTA + wid;
  const bool valid = env < P.n_envs;
  const uint32_t env_global = (uint32_t)(cfg.env_id_offset + env);
  const bool first_slice = P.set_budget >= 0 && slice == 0;
  int slice_left = P.slice_len > 0 ? P.slice_len : 0x7fffffff;

  int budget = 0;
  if (valid) budget = first_slice ? P.set_budget : P.budget[env];
  if (lane == 0) { S.pgs_on = 0; S.c.frozen = 0; S.c.parked = 0; S.cap_over = 0; S.cap_max = 0; }

  // ---------------------------------------------------------------- load the record (coalesced float4 -> smem staging -> registers)
  EnvRec* grec = P.rec + (valid ? env : 0);
  float pos[3], q[4], R[9], lvel[3], avel[3], facc[3], tacc[3];
  if (valid) {
    {
      const float4* src = reinterpret_cast<const float4*>(grec);
      float4* dst = reinterpret_cast<float4*>(&S.rowJ[0][0]);
      constexpr int N4 = sizeof(EnvRec) / 16;
      for (int i = lane; i < N4; i += 32) dst[i] = src[i];
    }
    __syncwarp();
    {
      const EnvRec* st = reinterpret_cast<const EnvRec*>(&S.rowJ[0][0]);
      const int b = lane < NB ? lane : 0;
#pragma unroll
      for (int k = 0; k < 3; k++) { pos[k] = st->body[F_POS + k][b]; lvel[k] = st->body[F_LVEL + k][b]; avel[k] = st->body[F_AVEL + k][b]; facc[k] = st->body[F_FACC + k][b]; tacc[k] = st->body[F_TACC + k][b]; }
#pragma unroll
      for (int k = 0; k < 4; k++) q[k] = st->body[F_QUAT + k][b];
      const uint32_t* cs = reinterpret_cast<const uint32_t*>(&st->c);
      uint32_t* cd = reinterpret_cast<uint32_t*>(&S.c);
      for (int i = lane; i < (int)(sizeof(Ctrl) / 4); i += 32) cd[i] = cs[i];
    }
    __syncwarp();
    if (first_slice && lane == 0) S.c.done_call = 0;
    quatToR(R, q);
    publish_body(S, lane, pos, q, R, lvel, avel);
    __syncwarp();
    joint_readouts(S, lane);  // derived from the pose: recomputed so that records written by set_state are consistent
    __syncwarp();
  } else {
#pragma unroll
    for (int k = 0; k < 3; k++) { pos[k] = 0; lvel[k] = 0; avel[k] = 0; facc[k] = 0; tacc[k] = 0; }
    q[0] = 1; q[1] = 0; q[2] = 0; q[3] = 0;
    quatToR(R, q);
  }

  const int nb_active = (cfg.scene == TFODE_SCENE_STEP_BOXES) ? NB : NBR;
  const int nj_active = (cfg.scene == TFODE_SCENE_STEP_BOXES) ? NJ : NJ - 2;
  const bool body_lane = lane < nb_active;
  const float h = cfg.dt;
  const float fps = 1.0f / h;
  const float invMass = body_lane ? M.invMass[lane] : 0.0f;
  const float ext_speed = (valid && P.ext_speed) ? P.ext_speed[env] : 0.0f;
  const int maxc = min(cfg.max_contacts, MAXC);
  float iI[6] = {0, 0, 0, 0, 0, 0};  // world inverse inertia of my body: xx xy xz yy yz zz

  // ================================================================== sub-step loop (CTA-uniform trip count)
  PT_DECL
  for (;;) {
    __syncwarp();
    if (valid && S.c.frozen) budget = 0;
    const bool runnable = valid && budget > 0 && slice_left > 0 && S.c.parked != 1;
    if (!__syncthreads_or(runnable ? 1 : 0)) break;
    const bool do_part_a = runnable && S.c.parked == 0;

    if (do_part_a) {
      // ------------------------------------------------------------ scene lambda prologue (rover_training_1.cc:129-136)
      if (lane == 0) {
        Ctrl& c = S.c;
        c.done_step = 0;
        if (P.ext_speed) c.robot_speed = fmin(fmax(-M.robot_max_speed, (double)ext_speed), M.robot_max_speed);
        else if (cfg.term > 0 && fabsf(c.speed) <= fabsf(cfg.speedf)) {
          c.speed += cfg.speedf / cfg.term * cfg.dt;
          c.robot_speed = fmin(fmax(-M.robot_max_speed, (double)c.speed), M.robot_max_speed);
        }
        float timef = (float)c.time_steps * cfg.dt;
        if (!c.ic_activated && cfg.ic_start >= 0 && (double)timef >= (double)c.ic_start) c.ic_activated = 1;
      }

      // ------------------------------------------------------------ collide (ode/environment.cc:56-150)
      int nc_total = 0;    // generated contacts (uniform)
      unsigned fork_hit = 0;
      unsigned hfpre[7] = {0, 0, 0, 0, 0, 0, 0};   // heightfield scene: spheres low enough to reach the map (one bit per sphere, sphere order)
      PC_DECL
      const int n_terrain = (cfg.scene == TFODE_SCENE_STEP_BOXES) ? 3 : (cfg.scene == TFODE_SCENE_HEIGHTFIELD ? 2 : 1);
      for (int t = 0; t < n_terrain; t++) {
        const int tkind = (t == 0) ? T_PLANE : (cfg.scene == TFODE_SCENE_STEP_BOXES ? T_BOX : T_HF);
        const int tbody = (tkind == T_BOX) ? NBR + (t - 1) : 15;
        const float* tpos = (tkind == T_BOX) ? S.bpos[tbody] : S.bpos[0];
        const float* tR = (tkind == T_BOX) ? S.bR[tbody] : S.bR[0];
        const float* tside = M.step_side[(tkind == T_BOX) ? (t - 1) : 0];
        unsigned box_wheels = 15u;   // wheels (bit w) whose tire can reach this step box
        if (tkind == T_BOX) {
          bool reach_w = false;
          if (lane < 4) {
            const float* wp = S.bpos[6 + lane];
            const float rad = 0.105f * 1.001f + 1e-4f;   // ring radius 0.055 + sphere radius 0.05 (ode/wheel.hh:63-73), grown by the margin
            reach_w = true;
#pragma unroll
            for (int k = 0; k < 3; k++) {
              const float range = 0.5f * (fabsf(tR[k * 3 + 0] * tside[0]) + fabsf(tR[k * 3 + 1] * tside[1]) + fabsf(tR[k * 3 + 2] * tside[2]));
              if (fabsf(wp[k] - tpos[k]) > range + rad) reach_w = false;
            }
          }
          box_wheels = __ballot_sync(FULLMASK, reach_w) & 15u;
        }
        for (int pass = 0; pass < 8; pass++) {
          CG cg[4]; int n = 0; int gbody = 0; int soft = 0;
          PC_MARK(3)
          if (pass == 0) {
            if (lane < NNS) {
              const int g = lane;
              gbody = M.ns_body[g];
              float gp[3], gR[9];
              const float* bR = S.bR[gbody];
              mulRv(gp, bR, M.ns_off[g]);
              gp[0] += S.bpos[gbody][0]; gp[1] += S.bpos[gbody][1]; gp[2] += S.bpos[gbody][2];
              if (M.ns_rot[g]) {  // body rotation x geom offset rotation (dGeomSetOffsetRotation, src/rover_1.cc:158-159)
                const float* oR = M.ns_offR[g];
#pragma unroll
                for (int r = 0; r < 3; r++)
#pragma unroll
                  for (int cc = 0; cc < 3; cc++) gR[r * 3 + cc] = tfn::dot3(bR[r * 3], bR[r * 3 + 1], bR[r * 3 + 2], oR[cc], oR[3 + cc], oR[6 + cc]);
              } else {
#pragma unroll
                for (int k = 0; k < 9; k++) gR[k] = bR[k];
              }
              const float* dim = M.ns_dim[g];
              // Conservative reject: the lowest point the geom can reach, against the highest point of the terrain piece
              // (0 for the plane, the top of the map, the top of the step box).  The margin is three orders of magnitude
              // above the rounding of these sums, so a skipped collider is one that would have returned no contact.
              const bool is_cyl = M.ns_type[g] == 2;
              const float reach = is_cyl ? (0.5f * dim[1] * fabsf(gR[8]) + dim[0])
                                         : 0.5f * (fabsf(gR[6]) * dim[0] + fabsf(gR[7]) * dim[1] + fabsf(gR[8]) * dim[2]);
              const float top = (tkind == T_PLANE) ? 0.0f
                              : (tkind == T_HF) ? (P.hf.pos[2] + P.hf.hmax)
                                                : (tpos[2] + 0.5f * (fabsf(tR[6]) * tside[0] + fabsf(tR[7]) * tside[1] + fabsf(tR[8]) * tside[2]));
              bool may_touch = !(gp[2] - reach > top + 1e-4f);
              // Step boxes: the geom's bounding sphere against the box's axis-aligned bounds.  Every sample point of the colliders below
              // (and the cylinder's own AABB of the fork callback) lies inside that sphere, so a geom that is clear of the bounds by the
              // same margin produces neither a contact nor a callback hit.
              bool far_box = false;
              if (tkind == T_BOX) {
                const float rad = is_cyl ? sqrtf(dim[0] * dim[0] + 0.25f * dim[1] * dim[1]) : 0.5f * sqrtf(dim[0] * dim[0] + dim[1] * dim[1] + dim[2] * dim[2]);
#pragma unroll
                for (int k = 0; k < 3; k++) {
                  const float range = 0.5f * (fabsf(tR[k * 3 + 0] * tside[0]) + fabsf(tR[k * 3 + 1] * tside[1]) + fabsf(tR[k * 3 + 2] * tside[2]));
                  if (fabsf(gp[k] - tpos[k]) > range + rad * 1.001f + 1e-4f) far_box = true;
                }
                if (far_box) may_touch = false;
              }
              if (is_cyl) {
                if (!may_touch) n = 0;
                else if (tkind == T_PLANE) n = col_cyl_plane(gp, gR, dim[0], dim[1], cg);
                else n = col_cyl_samples(tkind, gp, gR, dim[0], dim[1], tpos, tR, tside, P.hf, cg);
                if (M.ns_cb[g] && cfg.fork_callback && !far_box) {  // near-callback AABB test (Q10)
                  float lo[3], hi[3];
#pragma unroll
                  for (int k = 0; k < 3; k++) {
                    float rr = gR[k * 3 + 2];
                    float range = fabsf(rr * dim[1] * 0.5f) + dim[0] * sqrtf(fmaxf(0.0f, 1.0f - rr * rr));
                    lo[k] = gp[k] - range; hi[k] = gp[k] + range;
                  }
                  bool ov;
                  if (tkind == T_PLANE) ov = lo[2] <= 0.0f;
                  else if (tkind == T_BOX) {
                    ov = true;
#pragma unroll
                    for (int k = 0; k < 3; k++) {
                      float range = 0.5f * (fabsf(tR[k * 3 + 0] * tside[0]) + fabsf(tR[k * 3 + 1] * tside[1]) + fabsf(tR[k * 3 + 2] * tside[2]));
                      if (lo[k] > tpos[k] + range || hi[k] < tpos[k] - range) ov = false;
                    }
                  } else {
                    ov = !(lo[0] > P.hf.pos[0] + P.hf.length * 0.5f || hi[0] < P.hf.pos[0] - P.hf.length * 0.5f ||
                           lo[1] > P.hf.pos[1] + P.hf.width * 0.5f || hi[1] < P.hf.pos[1] - P.hf.width * 0.5f ||
                           lo[2] > P.hf.pos[2] + 1.0f || hi[2] < P.hf.pos[2] - 1.0f);
                  }
                  if (ov) fork_hit = 1;
                }
              } else {
                if (!may_touch) n = 0;
                else if (tkind == T_PLANE) n = col_box_plane(gp, gR, dim, cg);
                else n = col_box_samples(tkind, gp, gR, dim, tpos, tR, tside, P.hf, cg);
              }
            }
            if (tkind != T_PLANE) { PC_MARK(0) }
          } else if (tkind == T_HF) {
            // Candidate spheres were picked while the plane was tested (hfpre[]: bit = the sphere's lowest point is not
            // above the highest vertex of the map).  They are compacted in sphere order, 32 per round: a lane takes one
            // candidate through the block-maxima range test, then the whole warp scans the survivors one at a time.
            if (pass > 1) continue;
            int ncand = 0;
#pragma unroll
            for (int j = 0; j < 7; j++) ncand += __popc(hfpre[j]);
            for (int base_c = 0; base_c < ncand; base_c += 32) {
              int want = base_c + lane, s = -1;
#pragma unroll
              for (int j = 0; j < 7; j++) {
                const int pc = __popc(hfpre[j]);
                if (s < 0 && want >= 0 && want < pc) s = j * 32 + (int)__fns(hfpre[j], 0, want + 1);
                want -= pc;
              }
              float c[3] = {0, 0, 0}, cl[3] = {0, 0, 0}; int rng[4] = {0, 0, 0, 0};
              bool cand = false;
              if (s >= 0) {
                const int w = s / 50, k = s - w * 50;
                gbody = 6 + w; soft = (k == 49);
                const float4 rk = __ldg(P.ring4 + k);
                const float ring_k[3] = {rk.x, rk.y, rk.z};
                mulRv(c, S.bR[gbody], ring_k);
                c[0] += S.bpos[gbody][0]; c[1] += S.bpos[gbody][1]; c[2] += S.bpos[gbody][2];
                cand = hf_sphere_range(P.hf, c, M.sphere_r, cl, rng);
              }
              unsigned cm = __ballot_sync(FULLMASK, cand);
              PC_MARK(1)
              while (cm) {
                const int src = __ffs(cm) - 1; cm &= cm - 1;
                float cw_[3], cl_[3]; int rg_[4];
#pragma unroll
                for (int k = 0; k < 3; k++) { cw_[k] = __shfl_sync(FULLMASK, c[k], src); cl_[k] = __shfl_sync(FULLMASK, cl[k], src); }
#pragma unroll
                for (int k = 0; k < 4; k++) rg_[k] = __shfl_sync(FULLMASK, rng[k], src);
                const int info = __shfl_sync(FULLMASK, gbody | (tbody << 4) | (soft << 8), src);
                nc_total += col_sphere_hf_warp(P.hf, cw_, cl_, M.sphere_r, rg_, lane, S.cont, nc_total, maxc, info);
              }
              PC_MARK(2)
            }
            continue;
          } else {
            // Step boxes: a pass is skipped when none of the wheels its 32 spheres belong to can reach the box (wheel centre against
            // the box's axis-aligned bounds grown by the tire's outer radius; the skipped spheres would have returned no contact).
            if (tkind == T_BOX && !(box_wheels & (0x88C6231u >> (4 * (pass - 1)) & 15u))) continue;
            const int s = (pass - 1) * 32 + lane;
            float c[3] = {0, 0, 0};
            bool pre = false;
            if (s < NSPHERE) {
              const int w = s / 50, k = s - w * 50;
              gbody = 6 + w; soft = (k == 49);
              const float4 rk = __ldg(P.ring4 + k);
              const float ring_k[3] = {rk.x, rk.y, rk.z};
              mulRv(c, S.bR[gbody], ring_k);  // dBodyGetRelPointPos of the geom offset (ode/wheel.hh:70-72)
              c[0] += S.bpos[gbody][0]; c[1] += S.bpos[gbody][1]; c[2] += S.bpos[gbody][2];
              if (tkind == T_PLANE) n = col_sphere_plane(c, M.sphere_r, cg);
              else n = col_sphere_box(c, M.sphere_r, tpos, tR, tside, cg);
              if (cfg.scene == TFODE_SCENE_HEIGHTFIELD) pre = !((c[2] - P.hf.pos[2]) - M.sphere_r > P.hf.hmax);   // hf_to_local's height, hf_sphere_range's last test with zm <= hmax
            }
            if (tkind == T_PLANE && cfg.scene == TFODE_SCENE_HEIGHTFIELD) hfpre[pass - 1] = __ballot_sync(FULLMASK, pre);
          }
          if (__any_sync(FULLMASK, n > 0)) {
            int incl = warp_incl_scan(n, lane);
            int base = nc_total + incl - n;
            for (int i = 0; i < n; i++) {
              int slot = base + i;
              if (slot < maxc) {
                float* d = S.cont[slot];
                d[0] = cg[i].p[0]; d[1] = cg[i].p[1]; d[2] = cg[i].p[2]; d[3] = cg[i].n[0]; d[4] = cg[i].n[1]; d[5] = cg[i].n[2]; d[6] = cg[i].depth;
                d[7] = __int_as_float(gbody | (tbody << 4) | (soft << 8));
              }
            }
            nc_total += __shfl_sync(FULLMASK, incl, 31);
          }
        }
      }
      if (__any_sync(FULLMASK, fork_hit != 0) && lane == 0) S.c.fork_collision = 1;
      const int nc = min(nc_total, maxc);
      if (lane == 0) {
        S.c.n_contacts = nc_total;
        // capacity audit (the reference creates every contact, ode/environment.cc:101-128): tfode_batch_contact_stats reports these
        if (nc_total > maxc) S.cap_over++;
        if (nc_total > S.cap_max) S.cap_max = nc_total;
      }
      __syncwarp();

      PT_MARK(0)
      // ------------------------------------------------------------ islands: which rover groups touch which step box
      int lab_of_group[NISL];
      {
        unsigned adj3 = 0, adj4 = 0;
        if (lane < nc) {
          int info = __float_as_int(S.cont[lane][7]);
          int b1 = info & 15, b2 = (info >> 4) & 15;
          if (b2 == 10) adj3 = 1u << body_group(b1);
          if (b2 == 11) adj4 = 1u << body_group(b1);
        }
        adj3 = __reduce_or_sync(FULLMASK, adj3); adj4 = __reduce_or_sync(FULLMASK, adj4);
        unsigned S3 = (1u << 3) | adj3, S4 = (1u << 4) | adj4;
        if (S3 & S4) { S3 |= S4; S4 = S3; }
#pragma unroll
        for (int g = 0; g < NISL; g++) lab_of_group[g] = (S3 >> g & 1u) ? (__ffs(S3) - 1) : ((S4 >> g & 1u) ? (__ffs(S4) - 1) : g);
      }

      // ------------------------------------------------------------ per-body prep: world inverse inertia, gyroscopic torque, gravity, v/h + invM f
      if (body_lane) {
        const float i0 = M.invIb[lane][0], i1 = M.invIb[lane][1], i2 = M.invIb[lane][2];
        {
          int n = 0;
#pragma unroll
          for (int r = 0; r < 3; r++)
#pragma unroll
            for (int cc = r; cc < 3; cc++)
              iI[n++] = tfn::fma(R[r * 3 + 2], i2 * R[cc * 3 + 2], tfn::fma(R[r * 3 + 1], i1 * R[cc * 3 + 1], R[r * 3 + 0] * (i0 * R[cc * 3 + 0])));
        }
        // gyroscopic: tacc -= w x (I w), I w = R (Ib .* (R^T w))
        float wl[3]; mulRtv(wl, R, avel);
        wl[0] *= M.Ib[lane][0]; wl[1] *= M.Ib[lane][1]; wl[2] *= M.Ib[lane][2];
        float Iw[3]; mulRv(Iw, R, wl);
        float cr[3]; cross3(cr, avel, Iw);
        tacc[0] -= cr[0]; tacc[1] -= cr[1]; tacc[2] -= cr[2];
        facc[2] += M.mass[lane] * -9.80665f;
        float* bt = S.btmp[lane];
        bt[0] = facc[0] * invMass + lvel[0] * fps; bt[1] = facc[1] * invMass + lvel[1] * fps; bt[2] = facc[2] * invMass + lvel[2] * fps;
        { float t3[3]; symMulV(t3, iI, tacc);
          bt[3] = t3[0] + avel[0] * fps; bt[4] = t3[1] + avel[1] * fps; bt[5] = t3[2] + avel[2] * fps; }
        float* bi = S.binvI[lane];
#pragma unroll
        for (int k = 0; k < 6; k++) bi[k] = iI[k];
        bi[6] = invMass;
      }

      // ------------------------------------------------------------ joint rows (lane j = joint j)
      int jm = 0;        // rows of my joint
      int jlimit = 0; float jlimit_err = 0;
      float jfmax = 0, jvel = 0;
      if (lane < nj_active) {
        const int j = lane;
        if (!M.jslider[j]) {
          if (j == 1) { jfmax = 180.0f; jvel = S.c.servo_vel; }      // STEERING_MAX_TORQUE
          else if (j >= 3) { jfmax = S.c.wheel_fmax[j - 3]; jvel = S.c.wheel_vel[j - 3]; }
          jm = jfmax > 0 ? 6 : 5;
          if ((M.lostop[j] >= -3.14159265358979323846f || M.histop[j] <= 3.14159265358979323846f) && M.lostop[j] <= M.histop[j]) {
            float ang = S.c.jq_angle[j];
            if (ang <= M.lostop[j]) { jlimit = 1; jlimit_err = ang - M.lostop[j]; jm = 6; }
            else if (ang >= M.histop[j]) { jlimit = 2; jlimit_err = ang - M.histop[j]; jm = 6; }
          }
        } else jm = 6;  // lo = hi = 0: the limit row is always present
      }
      int jincl = warp_incl_scan(jm, lane);
      const int jrow0 = jincl - jm;
      const int njrows = __shfl_sync(FULLMASK, jincl, 31);
      const int m_total = njrows + 3 * nc;
      if (lane < nj_active) {
        const int j = lane;
        const int b1 = M.jb1[j], b2 = M.jb2[j];  // b2 = 15 : world
        const bool has2 = b2 < NB;
        const float* R1 = S.bR[b1]; const float* x1 = S.bpos[b1];
        const float* R2 = S.bR[has2 ? b2 : b1]; const float* x2 = S.bpos[has2 ? b2 : b1];
        const float k_erp = fps * cfg.erp;
        const int hdr = b1 | (b2 << 4) | (255 << 8);
        for (int r = 0; r < jm; r++) {
          float* Jr = S.rowJ[jrow0 + r];
#pragma unroll
          for (int k = 0; k < 12; k++) Jr[k] = 0.0f;
          const int Sr = jrow0 + r;
          ROWF(S, Sr, C) = 0.0f; ROWF(S, Sr, ADCFM) = cfg.cfm; ROWF(S, Sr, LO) = -CUDART_INF_F; ROWF(S, Sr, HI) = CUDART_INF_F; ROWF(S, Sr, LAM) = 0.0f;
          ROWF(S, Sr, HDR) = __int_as_float(hdr);
        }
        if (!M.jslider[j]) {
          // ball part (setBall)
          float a1[3], a2[3];
          mulRv(a1, R1, M.anchor1[j]); mulRv(a2, R2, M.anchor2[j]);
          for (int r = 0; r < 3; r++) {
            float* Jr = S.rowJ[jrow0 + r];
            Jr[r] = 1.0f; Jr[6 + r] = -1.0f;
            ROWF(S, jrow0 + r, C) = k_erp * (a2[r] + x2[r] - a1[r] - x1[r]);
          }
          { float* J0 = S.rowJ[jrow0 + 0]; float* J1 = S.rowJ[jrow0 + 1]; float* J2 = S.rowJ[jrow0 + 2];
            J0[4] = a1[2]; J0[5] = -a1[1]; J1[3] = -a1[2]; J1[5] = a1[0]; J2[3] = a1[1]; J2[4] = -a1[0];
            J0[10] = -a2[2]; J0[11] = a2[1]; J1[9] = a2[2]; J1[11] = -a2[0]; J2[9] = -a2[1]; J2[10] = a2[0]; }
          float ax1[3], ax2[3], pp[3], qq[3], bb[3];
          mulRv(ax1, R1, M.axis1[j]); mulRv(ax2, R2, M.axis2[j]);
          planeSpace(ax1, pp, qq);
          cross3(bb, ax1, ax2);
          { float* J3 = S.rowJ[jrow0 + 3]; float* J4 = S.rowJ[jrow0 + 4];
            for (int k = 0; k < 3; k++) { J3[3 + k] = pp[k]; J3[9 + k] = -pp[k]; J4[3 + k] = qq[k]; J4[9 + k] = -qq[k]; }
            ROWF(S, jrow0 + 3, C) = k_erp * dot3(bb, pp); ROWF(S, jrow0 + 4, C) = k_erp * dot3(bb, qq); }
          if (jm == 6) {
            float* J5 = S.rowJ[jrow0 + 5]; const int S5 = jrow0 + 5;
            for (int k = 0; k < 3; k++) { J5[3 + k] = ax1[k]; J5[9 + k] = -ax1[k]; }
            if (jlimit) {  // unpowered joint at a stop (the boggie hinge)
              ROWF(S, S5, C) = -k_erp * jlimit_err;
              if (jlimit == 1) { ROWF(S, S5, LO) = 0.0f; ROWF(S, S5, HI) = CUDART_INF_F; } else { ROWF(S, S5, LO) = -CUDART_INF_F; ROWF(S, S5, HI) = 0.0f; }
            } else { ROWF(S, S5, C) = jvel; ROWF(S, S5, LO) = -jfmax; ROWF(S, S5, HI) = jfmax; }
          }
        } else {
          // slider: battery clamp (two bodies) or step fix (body - world)
          const float* qrel = (j >= 7) ? S.c.fix_qrel[j - 7] : M.qrel[j];
          const float* offs = (j >= 7) ? S.c.fix_offset[j - 7] : M.offset[j];
          float qerr[4];
          if (has2) { float t4[4]; qmul_inv_a(t4, S.bq[b1], S.bq[b2]); qmul_inv_b(qerr, t4, qrel); }
          else qmul_inv_ab(qerr, S.bq[b1], qrel);
          if (qerr[0] < 0) { qerr[1] = -qerr[1]; qerr[2] = -qerr[2]; qerr[3] = -qerr[3]; }
          float e[3]; mulRv(e, R1, qerr + 1);
          for (int r = 0; r < 3; r++) {
            float* Jr = S.rowJ[jrow0 + r];
            Jr[3 + r] = 1.0f; if (has2) Jr[9 + r] = -1.0f;
            ROWF(S, jrow0 + r, C) = 2 * k_erp * e[r];
          }
          float ax1[3], pp[3], qq[3];
          mulRv(ax1, R1, M.axis1[j]);
          planeSpace(ax1, pp, qq);
          float* J3 = S.rowJ[jrow0 + 3]; float* J4 = S.rowJ[jrow0 + 4]; float* J5 = S.rowJ[jrow0 + 5];
          float cvec[3] = {0, 0, 0};
          float posn;
          if (has2) {
            for (int k = 0; k < 3; k++) cvec[k] = x2[k] - x1[k];
            float t3[3];
            cross3(t3, cvec, pp); for (int k = 0; k < 3; k++) { J3[3 + k] = 0.5f * t3[k]; J3[9 + k] = 0.5f * t3[k]; }
            cross3(t3, cvec, qq); for (int k = 0; k < 3; k++) { J4[3 + k] = 0.5f * t3[k]; J4[9 + k] = 0.5f * t3[k]; }
            for (int k = 0; k < 3; k++) { J3[6 + k] = -pp[k]; J4[6 + k] = -qq[k]; }
            float ofs[3]; mulRv(ofs, R2, offs);
            float cc[3] = {cvec[0] + ofs[0], cvec[1] + ofs[1], cvec[2] + ofs[2]};
            ROWF(S, jrow0 + 3, C) = k_erp * dot3(pp, cc); ROWF(S, jrow0 + 4, C) = k_erp * dot3(qq, cc);
            float qv[3] = {x1[0] - ofs[0] - x2[0], x1[1] - ofs[1] - x2[1], x1[2] - ofs[2] - x2[2]};
            posn = dot3(ax1, qv);
            // limit row with linear torque decoupling
            float cl[3] = {0.5f * cvec[0], 0.5f * cvec[1], 0.5f * cvec[2]}, ltd[3];
            cross3(ltd, cl, ax1);
            for (int k = 0; k < 3; k++) { J5[k] = ax1[k]; J5[6 + k] = -ax1[k]; J5[3 + k] = ltd[k]; J5[9 + k] = ltd[k]; }
          } else {
            float ofs[3] = {offs[0] - x1[0], offs[1] - x1[1], offs[2] - x1[2]};
            ROWF(S, jrow0 + 3, C) = k_erp * dot3(pp, ofs); ROWF(S, jrow0 + 4, C) = k_erp * dot3(qq, ofs);
            float qv[3] = {x1[0] - offs[0], x1[1] - offs[1], x1[2] - offs[2]};
            posn = dot3(ax1, qv);
            for (int k = 0; k < 3; k++) J5[k] = ax1[k];
          }
          for (int k = 0; k < 3; k++) { J3[k] = pp[k]; J4[k] = qq[k]; }
          ROWF(S, jrow0 + 5, C) = -k_erp * posn;  // lo = hi = 0: equality, unbounded
        }
      }
      // ------------------------------------------------------------ contact rows (lane c = contact c), dxJointContact::getInfo2
      if (lane < nc) {
        const float* cd = S.cont[lane];
        const int info = __float_as_int(cd[7]);
        const int b1 = info & 15, b2 = (info >> 4) & 15; const bool soft = (info >> 8) & 1;
        const bool has2 = b2 < NB;
        const int r0 = njrows + 3 * lane;
        float n[3] = {cd[3], cd[4], cd[5]};
        float c1[3] = {cd[0] - S.bpos[b1][0], cd[1] - S.bpos[b1][1], cd[2] - S.bpos[b1][2]};
        float c2[3] = {0, 0, 0};
        if (has2) { c2[0] = cd[0] - S.bpos[b2][0]; c2[1] = cd[1] - S.bpos[b2][1]; c2[2] = cd[2] - S.bpos[b2][2]; }
        float t1[3], t2[3]; planeSpace(n, t1, t2);
        const float erp_ = soft ? 0.4f : cfg.erp;
        float depth = cd[6]; if (depth < 0) depth = 0;
#pragma unroll
        for (int r = 0; r < 3; r++) {
          const float* d = (r == 0) ? n : (r == 1 ? t1 : t2);
          float* Jr = S.rowJ[r0 + r]; const int Sr = r0 + r;
          float x[3]; cross3(x, c1, d);
          Jr[0] = d[0]; Jr[1] = d[1]; Jr[2] = d[2]; Jr[3] = x[0]; Jr[4] = x[1]; Jr[5] = x[2];
          if (has2) { cross3(x, c2, d); Jr[6] = -d[0]; Jr[7] = -d[1]; Jr[8] = -d[2]; Jr[9] = -x[0]; Jr[10] = -x[1]; Jr[11] = -x[2]; }
          else { Jr[6] = 0; Jr[7] = 0; Jr[8] = 0; Jr[9] = 0; Jr[10] = 0; Jr[11] = 0; }
          ROWF(S, Sr, LAM) = 0.0f;
          if (r == 0) { ROWF(S, Sr, C) = fps * erp_ * depth; ROWF(S, Sr, ADCFM) = soft ? 0.005f : cfg.cfm; ROWF(S, Sr, LO) = 0.0f; ROWF(S, Sr, HI) = CUDART_INF_F; ROWF(S, Sr, HDR) = __int_as_float(b1 | (b2 << 4) | (255 << 8)); }
          else { ROWF(S, Sr, C) = 0.0f; ROWF(S, Sr, ADCFM) = cfg.cfm; ROWF(S, Sr, LO) = -cfg.mu; ROWF(S, Sr, HI) = cfg.mu; ROWF(S, Sr, HDR) = __int_as_float(b1 | (b2 << 4) | (r0 << 8)); }
        }
      }
      __syncwarp();

      // ------------------------------------------------------------ rhs, Ad (lane = row)
      for (int r = lane; r < m_total; r += 32) {
        const float* Jr = S.rowJ[r];
        const int hdr = __float_as_int(ROWF(S, r, HDR));
        const int b1 = hdr & 15, b2 = (hdr >> 4) & 15;
        const float* t1 = S.btmp[b1]; const float* i1 = S.binvI[b1];
        float s = tfn::dot6(Jr, t1);
        float a3[3]; symMulV(a3, i1, Jr + 3);
        float diag = tfn::fma(i1[6], dot3(Jr, Jr), dot3(a3, Jr + 3));
        if (b2 < NB) {
          const float* t2 = S.btmp[b2]; const float* i2 = S.binvI[b2];
          s = s + tfn::dot6(Jr + 6, t2);
          symMulV(a3, i2, Jr + 9);
          diag = diag + tfn::fma(i2[6], dot3(Jr + 6, Jr + 6), dot3(a3, Jr + 9));
        }
        const float cfm_h = ROWF(S, r, ADCFM) * fps;
        const float Ad = cfg.sor_w / (diag + cfm_h);
        ROWF(S, r, RHS) = tfn::fma(ROWF(S, r, C), fps, -s);
        ROWF(S, r, AD) = Ad;
        ROWF(S, r, ADCFM) = (SOLVER == TFODE_SOLVER_DANTZIG) ? cfm_h : Ad * cfm_h;   // the dense solver wants cfm / h itself on A's diagonal
        // c is consumed: its slot now holds the row's two pre-decoded sweep-queue entries (body 1's lane | body 2's lane << 16)
        ROWF(S, r, ENT) = __int_as_float(pgs_make_entry(r, hdr, b1) | ((b2 < NB ? pgs_make_entry(r, hdr, b2) : (b1 << 8)) << 16));
      }

      // ------------------------------------------------------------ constraint order per island (quickstep.cpp SOR_LCP)
      int cntN[NISL], cntF[NISL];
#pragma unroll
      for (int L = 0; L < NISL; L++) { cntN[L] = 0; cntF[L] = 0; }
      for (int r0 = 0; r0 < m_total; r0 += 32) {
        const int r = r0 + lane; const bool valid = r < m_total;
        int lab = -1, fric = 0;
        if (valid) { int hdr = __float_as_int(ROWF(S, r, HDR)); int g = body_group(hdr & 15); fric = ((hdr >> 8) & 255) != 255;
#pragma unroll
          for (int L = 0; L < NISL; L++) if (g == L) lab = lab_of_group[L]; }
#pragma unroll
        for (int L = 0; L < NISL; L++) {
          cntN[L] += __popc(__ballot_sync(FULLMASK, valid && lab == L && !fric));
          cntF[L] += __popc(__ballot_sync(FULLMASK, valid && lab == L && fric));
        }
      }
      int offL[NISL], mL[NISL], baseL[NISL];
      const int nshuf = (SOLVER == TFODE_SOLVER_DANTZIG) ? 0 : (cfg.pgs_iters + 7) >> 3;   // dWorldStep draws no random numbers
      { int acc = 0, draws = 0;
#pragma unroll
        for (int L = 0; L < NISL; L++) { mL[L] = cntN[L] + cntF[L]; offL[L] = acc; acc += mL[L]; baseL[L] = draws; draws += nshuf * max(mL[L] - 1, 0); }
        if (lane == 0) { S.n_draws = draws; S.c.n_rows = m_total; }
#pragma unroll
        for (int L = 0; L < NISL; L++) if (lane == L) { S.isl_off[L] = offL[L]; S.isl_m[L] = mL[L]; S.isl_base[L] = baseL[L]; S.isl_lab[L] = lab_of_group[L]; }
      }
      {
        int pN[NISL], pF[NISL];
#pragma unroll
        for (int L = 0; L < NISL; L++) { pN[L] = 0; pF[L] = 0; }
        const unsigned lt = (1u << lane) - 1u;
        for (int r0 = 0; r0 < m_total; r0 += 32) {
          const int r = r0 + lane; const bool valid = r < m_total;
          int lab = -1, fric = 0;
          if (valid) { int hdr = __float_as_int(ROWF(S, r, HDR)); int g = body_group(hdr & 15); fric = ((hdr >> 8) & 255) != 255;
#pragma unroll
            for (int L = 0; L < NISL; L++) if (g == L) lab = lab_of_group[L]; }
#pragma unroll
          for (int L = 0; L < NISL; L++) {
            unsigned bN = __ballot_sync(FULLMASK, valid && lab == L && !fric);
            unsigned bF = __ballot_sync(FULLMASK, valid && lab == L && fric);
            if (valid && lab == L) {
              int p = fric ? (offL[L] + mL[L] - 1 - (pF[L] + __popc(bF & lt))) : (offL[L] + pN[L] + __popc(bN & lt));
              S.order[p] = (unsigned char)r;
            }
            pN[L] += __popc(bN); pF[L] += __popc(bF);
          }
        }
      }
      __syncwarp();
      // Reorder 0 of every island (ODE shuffles before the first sweep).  The draws depend on the stream position only, so
      // all 32 lanes compute them at once (LCG jump-ahead to each draw's position) into the dead contact list; the swap
      // chain itself, the sequential part of Fisher-Yates, then runs with one leader lane per island.
      if constexpr (SOLVER != TFODE_SOLVER_DANTZIG) {
        unsigned char* dr = reinterpret_cast<unsigned char*>(&S.cont[0][0]);
        const uint32_t seed_now = S.c.lcg_seed;
#pragma unroll 1
        for (int L = 0; L < NISL; L++) {
          const int m_l = S.isl_m[L], off_l = S.isl_off[L], base_l = S.isl_base[L];
          for (int q = 1 + lane; q < m_l; q += 32) {
            uint32_t sd = lcg_jump(seed_now, (uint32_t)(base_l + q - 1));
            dr[off_l + q] = (unsigned char)lcg_randint_magic(sd, (uint32_t)(q + 1), c_fy_magic.v[q + 1]);
          }
        }
        __syncwarp();
        if (lane < NB) {
          const int g = body_group(lane), L = S.isl_lab[g];
          const int m_i = S.isl_m[L];
          if (lane == first_body_of_group(g) && L == g && m_i > 1) {
            unsigned char* o = S.order + S.isl_off[L]; const unsigned char* d = dr + S.isl_off[L];
            int sw = d[1];
            for (int q = 1; q < m_i; q++) {
              const int sw_next = d[q + 1 < m_i ? q + 1 : q];
              const unsigned char x = o[q], y = o[sw];
              o[q] = y; o[sw] = x;
              sw = sw_next;
            }
          }
        }
      }
      __syncwarp();

      if (lane < NB) {  // btmp becomes fc: the sweeps start from zero constraint acceleration
#pragma unroll
        for (int k = 0; k < 8; k++) S.btmp[lane][k] = 0.0f;
      }
      if (lane == 0) S.pgs_on = 1;
    } else if (lane == 0) S.pgs_on = 0;
    PT_MARK(1)
    __syncthreads();
    PT_MARK(2)

    // ---------------------------------------------------------------- PGS / SOR sweeps, packed over the CTA's environments
    if (SOLVER == TFODE_SOLVER_DANTZIG) { if (S.pgs_on) dantzig_env(S, P, env, lane); }
    else if (wid < PGS_WARPS) pgs_bodylane<LSU>(sm, cfg.pgs_iters, wid, lane);
    else if (cfg.pgs_iters > 8) pgs_helper(sm, cfg.pgs_iters, wid - PGS_WARPS, lane);   // permutations of the later reorders, off the sweeps' path
    PT_MARK(3)
    __syncthreads();
    PT_MARK(4)

    bool parked_now = false;
    if (do_part_a) {
      if (lane == 0) S.c.lcg_seed = lcg_jump(S.c.lcg_seed, (uint32_t)S.n_draws);
      // ------------------------------------------------------------ velocity update + dxStepBody
      if (body_lane) {
#pragma unroll
        const float* fc = S.btmp[lane];
#pragma unroll
        for (int k = 0; k < 3; k++) { lvel[k] += h * fc[k]; avel[k] += h * fc[3 + k]; }
        const float hm = h * invMass;
        float ht[3] = {tacc[0] * h, tacc[1] * h, tacc[2] * h};
#pragma unroll
        for (int k = 0; k < 3; k++) lvel[k] += hm * facc[k];
        { float t3[3]; symMulV(t3, iI, ht); avel[0] += t3[0]; avel[1] += t3[1]; avel[2] += t3[2]; }
#pragma unroll
        for (int k = 0; k < 3; k++) pos[k] += h * lvel[k];
        float dq[4];
        dq[0] = 0.5f * (-avel[0] * q[1] - avel[1] * q[2] - avel[2] * q[3]);
        dq[1] = 0.5f * (avel[0] * q[0] + avel[1] * q[3] - avel[2] * q[2]);
        dq[2] = 0.5f * (-avel[0] * q[3] + avel[1] * q[0] + avel[2] * q[1]);
        dq[3] = 0.5f * (avel[0] * q[2] - avel[1] * q[1] + avel[2] * q[0]);
#pragma unroll
        for (int k = 0; k < 4; k++) q[k] += h * dq[k];
        float l = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
        l = 1.0f / sqrtf(l);
#pragma unroll
        for (int k = 0; k < 4; k++) q[k] *= l;
        quatToR(R, q);
#pragma unroll
        for (int k = 0; k < 3; k++) { facc[k] = 0; tacc[k] = 0; }
      }
      publish_body(S, lane, pos, q, R, lvel, avel);
      __syncwarp();

      // ------------------------------------------------------------ Rover_1::next_step, first half (src/rover_1.cc:480-494)
      joint_readouts(S, lane);
      if (lane < 2) {  // FT_sensor::Update (ode/ft_sensor.cc:58-88); lane 0 front (main, front fork), lane 1 rear (boggie, rear fork)
        const int A = lane == 0 ? 0 : 3, B = lane == 0 ? 4 : 5;
        float rA[3], rB[3], pA[3], pB[3], vA[3], vB[3], t3[3];
        mulRv(rA, S.bR[A], M.ft_ocA[lane]); mulRv(rB, S.bR[B], M.ft_ocB[lane]);
        for (int k = 0; k < 3; k++) { pA[k] = rA[k] + S.bpos[A][k]; pB[k] = rB[k] + S.bpos[B][k]; }
        cross3(t3, &S.bvel[A][3], rA); for (int k = 0; k < 3; k++) vA[k] = S.bvel[A][k] + t3[k];
        cross3(t3, &S.bvel[B][3], rB); for (int k = 0; k < 3; k++) vB[k] = S.bvel[B][k] + t3[k];
        const float Kl[3] = {2e4f, 2e4f, 1e5f};
        float F[3], T[3];
        for (int k = 0; k < 3; k++) F[k] = Kl[k] * (pB[k] - pA[k]) + 5e2f * (vB[k] - vA[k]);
        const float* qA = S.bq[A]; const float* qB = S.bq[B];
        float n2 = qA[0] * qA[0] + qA[1] * qA[1] + qA[2] * qA[2] + qA[3] * qA[3];
        float qi[4] = {qA[0] / n2, -qA[1] / n2, -qA[2] / n2, -qA[3] / n2};
        float qr[4]; qmul(qr, qB, qi);
        float n = sqrtf(qr[1] * qr[1] + qr[2] * qr[2] + qr[3] * qr[3]);
        float angle, axis[3];
        if (n != 0.0f) { angle = 2.0f * tfn::atan2(n, fabsf(qr[0])); if (qr[0] < 0.0f) n = -n; axis[0] = qr[1] / n; axis[1] = qr[2] / n; axis[2] = qr[3] / n; }
        else { angle = 0; axis[0] = 1; axis[1] = 0; axis[2] = 0; }
        for (int k = 0; k < 3; k++) T[k] = angle * 1e3f * axis[k] + 1.0f * (S.bvel[B][3 + k] - S.bvel[A][3 + k]);
        for (int k = 0; k < 3; k++) { S.ftraw[lane * 6 + k] = F[k]; S.ftraw[lane * 6 + 3 + k] = T[k]; }
      }
      __syncwarp();
      if (lane == 0 || lane == 3 || lane == 4 || lane == 5) {  // dBodyAddForceAtRelPos / dBodyAddTorque on A (+) and B (-)
        const int s = (lane == 0 || lane == 4) ? 0 : 1;
        const bool isA = (lane == 0 || lane == 3);
        const float sg = isA ? 1.0f : -1.0f;
        const float* oc = isA ? M.ft_ocA[s] : M.ft_ocB[s];
        float F[3] = {sg * S.ftraw[s * 6 + 0], sg * S.ftraw[s * 6 + 1], sg * S.ftraw[s * 6 + 2]};
        float T[3] = {sg * S.ftraw[s * 6 + 3], sg * S.ftraw[s * 6 + 4], sg * S.ftraw[s * 6 + 5]};
        float pr[3]; mulRv(pr, R, oc);
        float cx[3]; cross3(cx, pr, F);
        for (int k = 0; k < 3; k++) { facc[k] += F[k]; tacc[k] += cx[k]; }
        for (int k = 0; k < 3; k++) tacc[k] += T[k];
      }
      if (lane < 12) {  // 12 biquads in place (Filters/cpp/filters.cc:137-144), fp64 state
        const double Cy1 = M.filt_c[0], Cy2 = M.filt_c[1], Cx0 = M.filt_c[2];
        double* f = S.c.filt[lane];
        const double x = (double)S.ftraw[lane];
        const double y = Cy1 * f[0] + Cy2 * f[1] + Cx0 * x + (2 * Cx0) * f[2] + Cx0 * f[3];
        f[1] = f[0]; f[0] = y; f[3] = f[2]; f[2] = x;
        S.c.ft[lane] = (float)y;
      }
      __syncwarp();
      if (lane == 0) {
        Ctrl& c = S.c;
        S.aux[5] = __int_as_float(0);
        c.ic_clock += (double)cfg.dt;
        if (c.ic_activated && c.ic_clock >= (double)c.ic_period) {
          const double delta_t = c.ic_clock;
          c.tick++;
          bool park = false;
          if (cfg.controller == TFODE_CTRL_RANDOM) {
            uint32_t r[4]; philox(cfg.seed, env_global, c.episode, c.tick, 1, r);
            c.steering_rate = fmin(fmax(-15.0, (double)u11(r[0]) * 15.0), 15.0);
            c.boggie_torque = fmin(fmax(-25.0, (double)u11(r[1]) * 25.0), 25.0);
          } else if (cfg.controller == TFODE_CTRL_MLP) {
            // _ComputeReward (src/rover_1_tf.cc:80-108)
            float np[3]; mulRv(np, S.bR[0], M.getpos_rel);
            np[0] += S.bpos[0][0]; np[1] += S.bpos[0][1]; np[2] += S.bpos[0][2];
            double reward = ((double)np[0] - c.last_pos[0]) * (25. / delta_t);
            reward -= fabs(c.boggie_torque) / 25.0 * 0.5;
            if (c.fork_collision) { reward -= 1; c.fork_collision = 0; }
            for (int k = 0; k < 3; k++) c.last_pos[k] = (double)np[k];
            c.total_reward += reward; c.last_reward = reward;
            float cur[17]; get_state17(S, cur);
            if (c.has_last_state) {
              if (c.pending_valid) append_transition(P, c, env_global, false, 0.0f);
              for (int i = 0; i < 17; i++) c.pending[i] = c.last_state[i];
              c.pending[17] = (float)c.steering_rate; c.pending[18] = (float)c.boggie_torque; c.pending[19] = (float)reward; c.pending_reward = reward;
              for (int i = 0; i < 17; i++) c.pending[20 + i] = cur[i];
              c.pending_valid = 1;
            }
            uint32_t r[4]; philox(cfg.seed, env_global, c.episode, c.tick, 1, r);
            const double draw = ((double)u11(r[0]) + 1) / 2;
            if (cfg.exploration && ((!c.explore && draw > 0.8) || (c.explore && draw > 0.7))) {
              c.explore = !c.explore;
              if (c.explore) { c.steering_rate = (double)u11(r[1]) * 15.0; c.boggie_torque = (double)u11(r[2]) * 25.0; }
            }
            if (!cfg.exploration || !c.explore) {
              const float scaling[17] = {90, 45, 25, 25, 45, 100, 100, 100, 30, 30, 30, 100, 100, 100, 30, 30, 30};
              if (cfg.policy_precision == TFODE_POLICY_FP32) {
                // fused actor: the environment's own warp runs TF_model::infer right here (actor_fused below), the two tanh
                // outputs become the servo / boggie setpoints before this sub-step's control block -- no hand-off, no second launch
                float* in = actor_scratch(S);
                for (int i = 0; i < 17; i++) in[i] = cur[i] / scaling[i];
                S.aux[5] = __int_as_float(1);
              } else {
                float* in = P.policy_in + (size_t)env * 17;
                for (int i = 0; i < 17; i++) in[i] = cur[i] / scaling[i];
                int slot = atomicAdd(P.n_parked, 1);
                P.parked_list[slot] = env;
                park = true;
              }
            }
            for (int i = 0; i < 17; i++) c.last_state[i] = cur[i];
            c.has_last_state = 1;
          } else if (cfg.controller == TFODE_CTRL_MODEL_TREE) {  // Rover_1_mt::_InternalControl, src/rover_1_mt.cc:72-83
            // GetState(flip = false, full = false) (:30-52): the 17-vector without the rear forces; the two trees are
            // walked by model_tree_kernel while the environment is parked and write _steering_rate/_boggie_torque.
            float cur[17]; get_state17(S, cur);
            float* in = P.policy_in + (size_t)env * 17;
            for (int i = 0; i < 11; i++) in[i] = cur[i];
            for (int i = 0; i < 3; i++) in[11 + i] = cur[14 + i];
            int slot = atomicAdd(P.n_parked, 1);
            P.parked_list[slot] = env;
            park = true;
          } else if (cfg.controller == TFODE_CTRL_CRAWLER) {  // src/rover_1_crawlers.cc:40-77
            const double ang = (double)(c.jq_angle[1] * 57.29577951308232f);
            switch (c.cr_phase) {
              case 1: c.boggie_torque += c.cr_dtorque_dt * delta_t; if (c.boggie_torque >= c.cr_amp) { c.boggie_torque = c.cr_amp; ++c.cr_phase; } break;
              case 2: c.steering_rate = c.cr_rate; if (ang >= 45.0) { c.steering_rate = 0; ++c.cr_phase; } break;
              case 3: c.boggie_torque -= c.cr_dtorque_dt * delta_t; if (c.boggie_torque <= -c.cr_amp) { c.boggie_torque = -c.cr_amp; ++c.cr_phase; } break;
              case 4: c.steering_rate = -c.cr_rate; if (ang <= -45.0) { c.steering_rate = 0; c.cr_phase = 1; } break;
            }
          }
          c.ic_clock = 0; c.ic_tick = 1;
          c.parked = park ? 1 : 0;
        } else c.ic_tick = 0;
      }
      __syncwarp();
      if (cfg.controller == TFODE_CTRL_MLP && __float_as_int(S.aux[5]) != 0) {   // warp-uniform: this environment's IC tick asks the actor
        actor_fused(S, P.policy_w, lane);
        if (lane == 0) {   // src/rover_1_tf.cc:184-185
          const float* y = actor_scratch(S);
          S.c.steering_rate = (double)y[0] * 15.0;
          S.c.boggie_torque = (double)y[1] * 25.0;
        }
        __syncwarp();
      }

      parked_now = S.c.parked == 1;  // hand the env to the policy kernel; part B runs in the next launch
    } else if (runnable) {
      // resumed after the policy kernel: TF_model::infer outputs are in policy_act (src/rover_1_tf.cc:184-185)
      if (lane == 0) {
        if (cfg.controller == TFODE_CTRL_MLP) {
          S.c.steering_rate = (double)S.c.policy_act[0] * 15.0;
          S.c.boggie_torque = (double)S.c.policy_act[1] * 25.0;
        }  // TFODE_CTRL_MODEL_TREE: model_tree_kernel wrote the two commands into the record
        S.c.parked = 0;
      }
      __syncwarp();

    }

    if (runnable && !parked_now) {
    // ---------------------------------------------------------------- Rover_1::next_step, second half (src/rover_1.cc:496-502)
    if (lane == 0) {
      Ctrl& c = S.c;
      // _UpdateWheelControl (:346-386)
      const float gamma = c.jq_angle[1], dgamma = c.jq_rate[1];
      const float wb = 0.58f, wt = 0.61f, rs = (float)c.robot_speed, wr = 0.105f, WMS = 7.4351f;
      float diff[4], trans[4], min_speed = 0, adj;
      const float tg = tfn::tan(gamma / 2);
      const bool crawl = cfg.crawling_mode || cfg.controller == TFODE_CTRL_CRAWLER;
#pragma unroll
      for (int i = 0; i < 4; i++) {
        diff[i] = (i % 2 ? -1 : 1) * wt / wb * tg;
        trans[i] = (i / 2 ? -1 : 1) * (-wb * tg + (i % 2 ? -1 : 1) * wt) * dgamma / 4;
        if (crawl) { if (rs >= 0) min_speed = fmaxf(min_speed, -trans[i] / (1 + diff[i])); else min_speed = fminf(min_speed, -trans[i] / (1 + diff[i])); }
      }
      if (rs >= 0) adj = fmaxf(rs, min_speed); else adj = fminf(rs, min_speed);
#pragma unroll
      for (int i = 0; i < 4; i++) { if (rs >= 0) adj = fminf(adj, (WMS * wr - trans[i]) / (1 + diff[i])); else adj = fmaxf(adj, (-WMS * wr - trans[i]) / (1 + diff[i])); }
#pragma unroll
      for (int i = 0; i < 4; i++) {
        float W = (adj * (1 + diff[i]) + trans[i]) / wr;
        W = fminf(fmaxf(-WMS, W), WMS);  // _ApplyWheelControl (:389-400)
        c.W[i] = W;
        float mt = 25.4f - 3.416228430014391f * fabsf(c.jq_rate[3 + i]);
        if (mt >= 0) c.wheel_fmax[i] = mt;  // dParamFMax ignores negative values
        c.wheel_vel[i] = W;
      }
      // Servo::set_desired_vel (ode/servo.cc:123-134) + Servo::next_step (:145-167)
      const float vmax = (float)(15 * 0.017453292519943295), smin = (float)(-45 * 0.017453292519943295), smax = (float)(45 * 0.017453292519943295);
      float vel = (float)(c.steering_rate * 0.017453292519943295);
      vel = fmaxf(-vmax, vel); vel = fminf(vel, vmax);
      if ((gamma <= smin && vel < 0) || (gamma >= smax && vel > 0)) vel = 0;
      c.servo_vel = vel; c.servo_mode_vel = 1;
      // _ApplyBoggieControl (:411-416): torque vector on the boggie (+) and the rear body (-)
      const double bmax = (cfg.controller == TFODE_CTRL_CRAWLER) ? (double)c.cr_amp : 25.0;
      const float tq = (float)fmin(fmax(-bmax, c.boggie_torque), bmax);
      float ax[3]; mulRv(ax, S.bR[3], M.axis1[2]);
      S.aux[0] = ax[0] * tq; S.aux[1] = ax[1] * tq; S.aux[2] = ax[2] * tq;
    }
    __syncwarp();
    if (lane == 3) { tacc[0] += S.aux[0]; tacc[1] += S.aux[1]; tacc[2] += S.aux[2]; }
    if (lane == 2) { tacc[0] += -S.aux[0]; tacc[1] += -S.aux[1]; tacc[2] += -S.aux[2]; }

    // ---------------------------------------------------------------- termination test (rover_training_1.cc:141-143) + epilogue (:212-224)
    if (lane == 0) {
      Ctrl& c = S.c;
      int reason = 0;
      if (cfg.terminate) {
        float p[3]; mulRv(p, S.bR[0], M.getpos_rel);
        p[0] += S.bpos[0][0]; p[1] += S.bpos[0][1]; p[2] += S.bpos[0][2];
        const double time = (double)((float)c.time_steps * cfg.dt);
        const bool upside = S.bR[0][8] < 0;
        if (time >= (double)cfg.timeout) reason = TFODE_TERM_TIMEOUT;
        else if (cfg.y_max > 0 && fabs((double)p[1]) >= (double)cfg.y_max) reason = TFODE_TERM_Y_MAX;
        else if ((cfg.x_goal_abs ? fabs((double)p[0]) : (double)p[0]) >= (double)cfg.x_goal) reason = TFODE_TERM_X_GOAL;
        else if (cfg.check_upside_down && upside) reason = TFODE_TERM_UPSIDE_DOWN;
        if (reason) {
          if (c.pending_valid) {
            float pen = 0;
            if (cfg.y_max > 0 && fabs((double)p[1]) >= (double)cfg.y_max) pen = 2; else if (cfg.check_upside_down && upside) pen = 5;
            append_transition(P, c, env_global, true, pen); c.pending_valid = 0;
          }
          c.last_ep_return = (float)c.total_reward; c.last_ep_x = p[0]; c.last_ep_y = p[1]; c.last_ep_steps = (int)c.time_steps;
          c.done_call = 1; c.done_step = 1; c.term_reason = reason; c.episodes_done++;
          if (!cfg.auto_reset) c.frozen = 1;
        }
      }
      if (!reason) c.time_steps++;
      S.aux[4] = __int_as_float(reason);
    }
    __syncwarp();

    budget--; slice_left--;
    if (__float_as_int(S.aux[4]) != 0 && cfg.auto_reset) {
      const uint32_t ep = S.c.episode + 1;
      __syncwarp();
      reset_env(S, P, env_global, ep, lane, pos, q, lvel, avel, facc, tacc);
      quatToR(R, q);
      publish_body(S, lane, pos, q, R, lvel, avel);
      __syncwarp();
      joint_readouts(S, lane);
      __syncwarp();
    }

    }
    PT_MARK(5)
  }

  PT_FLUSH
  // ---------------------------------------------------------------- store the record
  __syncwarp();
  if (valid) {
  if (lane == 0) {
    P.budget[env] = budget;
    if (S.cap_over) atomicAdd(P.cstats, (unsigned long long)S.cap_over);
    if (S.cap_max > 0) atomicMax(P.cstats + 1, (unsigned long long)S.cap_max);
  }
  {
    EnvRec* st = reinterpret_cast<EnvRec*>(&S.rowJ[0][0]);
    if (lane < NB) {
#pragma unroll
      for (int k = 0; k < 3; k++) { st->body[F_POS + k][lane] = pos[k]; st->body[F_LVEL + k][lane] = lvel[k]; st->body[F_AVEL + k][lane] = avel[k]; st->body[F_FACC + k][lane] = facc[k]; st->body[F_TACC + k][lane] = tacc[k]; }
#pragma unroll
      for (int k = 0; k < 4; k++) st->body[F_QUAT + k][lane] = q[k];
    }
    const uint32_t* cs = reinterpret_cast<const uint32_t*>(&S.c);
    uint32_t* cd = reinterpret_cast<uint32_t*>(&st->c);
    for (int i = lane; i < (int)(sizeof(Ctrl) / 4); i += 32) cd[i] = cs[i];
    __syncwarp();
    const float4* src = reinterpret_cast<const float4*>(st);
    float4* dst = reinterpret_cast<float4*>(grec);
    constexpr int N4 = sizeof(EnvRec) / 16;
    for (int i = lane; i < N4; i += 32) dst[i] = src[i];
  }
  }
  if (P.slice_len <= 0) return;
  __threadfence();     // this item's records are visible device-wide ...
  __syncthreads();
  if (threadIdx.x == 0) st_release(P.slice_done + group, slice + 1);   // ... before the group's next slice may be claimed
  }
}


// ------------------------------------------------------------------------------------------------
// reset kernel: construction of env records (initial create, tfode_batch_reset)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(WARPS_PER_CTA * 32) reset_kernel(const StepParams P, const int* env_ids, int n_ids, int first_time) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int idx = blockIdx.x * WARPS_PER_CTA + wid;
  if (idx >= n_ids) return;
  const int env = env_ids ? env_ids[idx] : idx;
  if (env < 0 || env >= P.n_envs) return;
  WarpSm& S = reinterpret_cast<WarpSm*>(smem_raw)[wid];
  EnvRec* grec = P.rec + env;
  {
    const uint32_t* cs = reinterpret_cast<const uint32_t*>(&grec->c);
    uint32_t* cd = reinterpret_cast<uint32_t*>(&S.c);
    for (int i = lane; i < (int)(sizeof(Ctrl) / 4); i += 32) cd[i] = first_time ? 0u : cs[i];
  }
  __syncwarp();
  const uint32_t ep = first_time ? 0u : S.c.episode + 1;
  if (lane == 0) { S.c.done_call = 0; S.c.frozen = 0; }
  float pos[3], q[4], R[9], lvel[3], avel[3], facc[3], tacc[3];
  reset_env(S, P, (uint32_t)(P.cfg.env_id_offset + env), ep, lane, pos, q, lvel, avel, facc, tacc);
  if (lane == 0) { S.c.frozen = 0; S.c.parked = 0; }
  quatToR(R, q);
  publish_body(S, lane, pos, q, R, lvel, avel);
  __syncwarp();
  joint_readouts(S, lane);
  __syncwarp();
  if (lane < NB) {
    for (int k = 0; k < 3; k++) { grec->body[F_POS + k][lane] = pos[k]; grec->body[F_LVEL + k][lane] = 0; grec->body[F_AVEL + k][lane] = 0; grec->body[F_FACC + k][lane] = 0; grec->body[F_TACC + k][lane] = 0; }
    for (int k = 0; k < 4; k++) grec->body[F_QUAT + k][lane] = q[k];
  }
  {
    const uint32_t* cs = reinterpret_cast<const uint32_t*>(&S.c);
    uint32_t* cd = reinterpret_cast<uint32_t*>(&grec->c);
    for (int i = lane; i < (int)(sizeof(Ctrl) / 4); i += 32) cd[i] = cs[i];
  }
  if (lane == 0) P.budget[env] = 0;
}

// observation / bodies gathers (thread per env)
__global__ void observe_kernel(const EnvRec* rec, int n, float* obs17, float* reward, uint8_t* done, int32_t* ncont) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  const EnvRec& r = rec[e];
  if (obs17) {
    float q[4] = {r.body[F_QUAT + 0][0], r.body[F_QUAT + 1][0], r.body[F_QUAT + 2][0], r.body[F_QUAT + 3][0]};
    float R[9]; quatToR(R, q);
    float vx0 = R[0], vx1 = R[3], vx2 = R[6];
    float direction = tfn::asin(vx1) * 57.29577951308232f;
    if (vx0 < 0) direction = (vx1 > 0 ? 1.0f : -1.0f) * 180.0f - direction;
    float* o = obs17 + (size_t)e * 17;
    o[0] = direction; o[1] = r.c.jq_angle[1] * 57.29577951308232f; o[2] = tfn::asin(R[7]) * 57.29577951308232f;
    o[3] = tfn::asin(-vx2) * 57.29577951308232f; o[4] = r.c.jq_angle[2] * 57.29577951308232f;
    for (int i = 0; i < 12; i++) o[5 + i] = r.c.ft[i];
  }
  if (reward) reward[e] = (float)r.c.last_reward;
  if (done) done[e] = (uint8_t)(r.c.done_call != 0);
  if (ncont) ncont[e] = r.c.n_contacts;
}
// Rollout rows for the trainer (K7): what one policy tick hands back per environment, packed so that ONE collective moves it --
// obs17 (Rover_1_tf::GetState), reward of the last IC tick, done flag of the last call, return and length of the last finished
// episode (the `[Success] ... Rmoy` line of src/rover_training_1.cc:203-209), global environment id.
__global__ void pack_rollout_kernel(const EnvRec* rec, int n, int env_id_offset, float* rows) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  const EnvRec& r = rec[e];
  float q[4] = {r.body[F_QUAT + 0][0], r.body[F_QUAT + 1][0], r.body[F_QUAT + 2][0], r.body[F_QUAT + 3][0]};
  float R[9]; quatToR(R, q);
  const float vx0 = R[0], vx1 = R[3], vx2 = R[6];
  float direction = tfn::asin(vx1) * 57.29577951308232f;
  if (vx0 < 0) direction = (vx1 > 0 ? 1.0f : -1.0f) * 180.0f - direction;
  float* o = rows + (size_t)e * TFODE_ROLLOUT_WIDTH;
  o[0] = direction; o[1] = r.c.jq_angle[1] * 57.29577951308232f; o[2] = tfn::asin(R[7]) * 57.29577951308232f;
  o[3] = tfn::asin(-vx2) * 57.29577951308232f; o[4] = r.c.jq_angle[2] * 57.29577951308232f;
  for (int i = 0; i < 12; i++) o[5 + i] = r.c.ft[i];
  o[17] = (float)r.c.last_reward; o[18] = r.c.done_call != 0 ? 1.0f : 0.0f;
  o[19] = r.c.last_ep_return; o[20] = (float)r.c.last_ep_steps; o[21] = (float)r.c.episodes_done;
  o[22] = (float)(env_id_offset + e); o[23] = 0.0f;
}
__global__ void bodies_kernel(const EnvRec* rec, int n, float* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * NB) return;
  const int e = i / NB, b = i - e * NB;
  const EnvRec& r = rec[e];
  float* o = out + (size_t)i * 13;
  for (int k = 0; k < 13; k++) o[k] = r.body[k][b];
}
__global__ void set_actions_kernel(EnvRec* rec, int n, const float* act, int crawler) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  // boggie_max_torque is 25 (src/rover_1.cc:50) except for Crawler_1, whose constructor sets it to ITS torque amplitude
  // (src/rover_1_crawlers.cc:22) -- per environment here, since BASELINE config 4 jitters the amplitude
  const double bmax = crawler ? (double)rec[e].c.cr_amp : 25.0;
  rec[e].c.steering_rate = fmin(fmax(-15.0, (double)act[e * 2]), 15.0);          // SetSteeringRate, src/rover_1.cc:326-329
  rec[e].c.boggie_torque = fmin(fmax(-bmax, (double)act[e * 2 + 1]), bmax);  // SetBoggieTorque :338-341
}

// ------------------------------------------------------------------------------------------------
// fp32 policy kernel (TF_model<float>::infer for the parked envs): 17 -> 512 relu -> 512 relu -> 2 tanh
// ------------------------------------------------------------------------------------------------
constexpr int POL_ROWS = 8;
__global__ void __launch_bounds__(256) policy_fp32_kernel(const float* __restrict__ in_all, const int* __restrict__ list, const int* __restrict__ n_list, int m_direct,
                                                          const float* __restrict__ W1, const float* __restrict__ b1, const float* __restrict__ W2, const float* __restrict__ b2,
                                                          const float* __restrict__ W3, const float* __restrict__ b3, EnvRec* rec, float* act_direct) {
  __shared__ float x[POL_ROWS][17];
  __shared__ float h1[POL_ROWS][512];
  __shared__ float h2[POL_ROWS][512];
  const int m = list ? *n_list : m_direct;
  const int row0 = blockIdx.x * POL_ROWS;
  if (row0 >= m) return;
  const int nr = min(POL_ROWS, m - row0);
  const int t = threadIdx.x;
  for (int i = t; i < nr * 17; i += 256) { int r = i / 17, k = i - r * 17; int src = list ? list[row0 + r] : (row0 + r); x[r][k] = in_all[(size_t)src * 17 + k]; }
  __syncthreads();
  for (int j = t; j < 512; j += 256) {
    float acc[POL_ROWS];
#pragma unroll
    for (int r = 0; r < POL_ROWS; r++) acc[r] = b1[j];
    for (int i = 0; i < 17; i++) { float w = W1[i * 512 + j];
#pragma unroll
      for (int r = 0; r < POL_ROWS; r++) acc[r] = tfn::fma(x[r][i], w, acc[r]); }
#pragma unroll
    for (int r = 0; r < POL_ROWS; r++) h1[r][j] = acc[r] > 0 ? acc[r] : 0;
  }
  __syncthreads();
  for (int j = t; j < 512; j += 256) {
    float acc[POL_ROWS];
#pragma unroll
    for (int r = 0; r < POL_ROWS; r++) acc[r] = b2[j];
    for (int i = 0; i < 512; i++) { float w = W2[i * 512 + j];
#pragma unroll
      for (int r = 0; r < POL_ROWS; r++) acc[r] = tfn::fma(h1[r][i], w, acc[r]); }
#pragma unroll
    for (int r = 0; r < POL_ROWS; r++) h2[r][j] = acc[r] > 0 ? acc[r] : 0;
  }
  __syncthreads();
  // layer 3: one thread per (row, output), accumulated in input order from the bias like the two layers above
  // (the fp32 policy path is bit-reproducible; the tensor-core path is the fast one)
  if (t < nr * 2) {
    const int r = t >> 1, j = t & 1;
    float s3 = b3[j];
    for (int i = 0; i < 512; i++) s3 = tfn::fma(h2[r][i], W3[i * 2 + j], s3);
    const float a = tfn::tanh(s3);
    if (list) { EnvRec& e = rec[list[row0 + r]]; e.c.policy_act[j] = a; if (j == 0) e.c.parked = 2; }
    else act_direct[(size_t)(row0 + r) * 2 + j] = a;
  }
}
__global__ void zero_int_kernel(int* p) { *p = 0; }

}  // namespace tfode
