// The fused per-environment step kernel (sm_100a): one warp = one rover environment,
// lane b (< 12) = rigid body b.  Everything one call of the reference's step lambda does
// (src/rover_training_1.cc:127-146 -> ode/environment.hh:127-136 -> src/rover_1.cc:478-503)
// runs inside this kernel for S sub-steps with the environment resident in registers and
// shared memory; HBM is touched once per launch (record load/store).
//
//   phase            lanes work on                    shared memory
//   collide          geoms (11 non-sphere, 200 tire spheres), warp scans  contact list
//   rows             joints (9) / contacts (<=24)     per-env Jacobian rows J[126][12] + row scalars
//   rhs / Ad         rows                             -
//   order            rows (ballots), island leaders (Fisher-Yates with ODE's LCG)
//   PGS sweep        bodies: the two lanes of a row's bodies form the 12-wide dot product
//                    with one shuffle; islands advance concurrently (they share no body)
//   integrate        bodies
//   robot.next_step  FT springs (2 lanes), biquads (12 lanes, fp64), control scalars (lane 0)
#pragma once
#include <math_constants.h>

#include "../../include/tfode_numerics.h"
#include "tfode_device.cuh"

namespace tfode {

__constant__ Model g_model;

#define FULLMASK 0xffffffffu

struct __align__(16) WarpSm {
  float bpos[NB][4];
  float bq[NB][4];
  float bR[NB][12];
  float bvel[NB][8];
  float binvI[NB][8];   // xx xy xz yy yz zz invMass -
  float btmp[NB][8];    // v/h + invM*f while the right-hand sides are formed; then fc, the constraint acceleration
                        // M^-1 J^T lambda accumulated by the sweeps (6 of 8 used)
  float cont[MAXC][8];  // pos, normal, depth, info bits (b1 | b2<<4 | soft<<8)
  float rowJ[MAXROWS][12];
  // Row scalars.  Default layout: split by access pattern so that the sweeps' per-lane gathers (one row per lane, any row)
  // spread over the banks -- {rhs, Ad, Ad*cfm, lo} as one 16-byte load (bank quad = row mod 8), {hi, lambda} as one 8-byte
  // load (bank pair = row mod 16), header and c/entry words on their own.  The former 32-byte row record put every row's
  // lambda on 4 of the 32 banks (8*row + 5) and the float4 on 4 of the 8 bank quads.  -DTFODE_ROWS_AOS restores it (A/B).
#ifdef TFODE_ROWS_AOS
  float rowS[MAXROWS][8];  // rhs, Ad, Ad*cfm, lo, hi, lambda, hdr, (c during build)
#else
  float4 rowC[MAXROWS];    // rhs, Ad, Ad*cfm, lo
  float2 rowHL[MAXROWS];   // hi, lambda
  float rowHdr[MAXROWS];   // body1 | body2 << 4 | friction-normal row << 8 (bytes 0/1 later: the order arrays of reorders 1 and 2)
  float rowCE[MAXROWS];    // c while the rows are built, then the row's two pre-decoded sweep-queue entries
#endif
  unsigned char order[MAXROWS + 2];
  int isl_off[NISL + 1], isl_m[NISL + 1], isl_base[NISL + 1];
  int isl_lab[NISL + 1];  // island label of each base group (a group merged with a step box takes the smallest group index)
  int pgs_on, n_draws;
  int cap_over, cap_max;   // contact-capacity audit of this launch: sub-steps over capacity, largest generated count
  float ftraw[12];
  float aux[12];
  Ctrl c;
  float bank_skew[4];   // sizeof % 128 == 16: the same row of neighbouring environments starts 4 banks apart
};
static_assert(sizeof(EnvRec) <= sizeof(float) * MAXROWS * 12, "record staging aliases the row array");

enum { RS_RHS = 0, RS_AD = 1, RS_ADCFM = 2, RS_LO = 3, RS_HI = 4, RS_LAM = 5, RS_HDR = 6, RS_C = 7, RS_ENT = 7 };
// lvalue of scalar `f` (an RS_* name without the prefix) of row r
#ifdef TFODE_ROWS_AOS
#define ROWF(S, r, f) ((S).rowS[(r)][RS_##f])
#define ROW_HDR_STRIDE 32
#else
#define ROWF_RHS(S, r) ((S).rowC[(r)].x)
#define ROWF_AD(S, r) ((S).rowC[(r)].y)
#define ROWF_ADCFM(S, r) ((S).rowC[(r)].z)
#define ROWF_LO(S, r) ((S).rowC[(r)].w)
#define ROWF_HI(S, r) ((S).rowHL[(r)].x)
#define ROWF_LAM(S, r) ((S).rowHL[(r)].y)
#define ROWF_HDR(S, r) ((S).rowHdr[(r)])
#define ROWF_C(S, r) ((S).rowCE[(r)])
#define ROWF_ENT(S, r) ((S).rowCE[(r)])
#define ROWF(S, r, f) ROWF_##f(S, r)
#define ROW_HDR_STRIDE 4
#endif

// ---------------------------------------------------------------- device-scope flags of the persistent work queue
__device__ __forceinline__ int ld_acquire(const int* p) { int v; asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void st_release(int* p, int v) { asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }

// ---------------------------------------------------------------- small math
// fma placement of these products is fixed by the numerics contract (include/tfode_numerics.h); the kernels are
// compiled with -fmad=false, so every other expression below is evaluated exactly as written.
__device__ __forceinline__ float dot3(const float* a, const float* b) { return tfn::dot3(a, b); }
__device__ __forceinline__ void cross3(float* r, const float* a, const float* b) { tfn::cross3(r, a, b); }
__device__ __forceinline__ void mulRv(float* r, const float* R, const float* v) { tfn::mulRv(r, R, v, 3); }    // R: 3 rows of stride 3
__device__ __forceinline__ void mulRtv(float* r, const float* R, const float* v) { tfn::mulRtv(r, R, v, 3); }
__device__ __forceinline__ void symMulV(float* r, const float* S, const float* v) {  // S = xx xy xz yy yz zz
  float a = tfn::dot3(S[0], S[1], S[2], v[0], v[1], v[2]);
  float b = tfn::dot3(S[1], S[3], S[4], v[0], v[1], v[2]);
  float c = tfn::dot3(S[2], S[4], S[5], v[0], v[1], v[2]);
  r[0] = a; r[1] = b; r[2] = c;
}
__device__ __forceinline__ void quatToR(float* R, const float* q) {
  float qq1 = 2 * q[1] * q[1], qq2 = 2 * q[2] * q[2], qq3 = 2 * q[3] * q[3];
  R[0] = 1 - qq2 - qq3; R[1] = 2 * (q[1] * q[2] - q[0] * q[3]); R[2] = 2 * (q[1] * q[3] + q[0] * q[2]);
  R[3] = 2 * (q[1] * q[2] + q[0] * q[3]); R[4] = 1 - qq1 - qq3; R[5] = 2 * (q[2] * q[3] - q[0] * q[1]);
  R[6] = 2 * (q[1] * q[3] - q[0] * q[2]); R[7] = 2 * (q[2] * q[3] + q[0] * q[1]); R[8] = 1 - qq1 - qq2;
}
__device__ __forceinline__ void planeSpace(const float* n, float* p, float* q) {
  if (fabsf(n[2]) > 0.70710678118654752440f) {
    float a = n[1] * n[1] + n[2] * n[2];
    float k = 1.0f / sqrtf(a);
    p[0] = 0; p[1] = -n[2] * k; p[2] = n[1] * k;
    q[0] = a * k; q[1] = -n[0] * p[2]; q[2] = n[0] * p[1];
  } else {
    float a = n[0] * n[0] + n[1] * n[1];
    float k = 1.0f / sqrtf(a);
    p[0] = -n[1] * k; p[1] = n[0] * k; p[2] = 0;
    q[0] = -n[2] * p[1]; q[1] = n[2] * p[0]; q[2] = a * k;
  }
}
__device__ __forceinline__ void qmul_inv_a(float* r, const float* b, const float* c) {  // b^-1 * c
  r[0] = b[0] * c[0] + b[1] * c[1] + b[2] * c[2] + b[3] * c[3];
  r[1] = b[0] * c[1] - b[1] * c[0] - b[2] * c[3] + b[3] * c[2];
  r[2] = b[0] * c[2] - b[2] * c[0] - b[3] * c[1] + b[1] * c[3];
  r[3] = b[0] * c[3] - b[3] * c[0] - b[1] * c[2] + b[2] * c[1];
}
__device__ __forceinline__ void qmul_inv_b(float* r, const float* b, const float* c) {  // b * c^-1
  r[0] = b[0] * c[0] + b[1] * c[1] + b[2] * c[2] + b[3] * c[3];
  r[1] = -b[0] * c[1] + b[1] * c[0] - b[2] * c[3] + b[3] * c[2];
  r[2] = -b[0] * c[2] + b[2] * c[0] - b[3] * c[1] + b[1] * c[3];
  r[3] = -b[0] * c[3] + b[3] * c[0] - b[1] * c[2] + b[2] * c[1];
}
__device__ __forceinline__ void qmul_inv_ab(float* r, const float* b, const float* c) {  // b^-1 * c^-1
  r[0] = b[0] * c[0] - b[1] * c[1] - b[2] * c[2] - b[3] * c[3];
  r[1] = -b[0] * c[1] - b[1] * c[0] + b[2] * c[3] - b[3] * c[2];
  r[2] = -b[0] * c[2] - b[2] * c[0] + b[3] * c[1] - b[1] * c[3];
  r[3] = -b[0] * c[3] - b[3] * c[0] + b[1] * c[2] - b[2] * c[1];
}
__device__ __forceinline__ void qmul(float* r, const float* b, const float* c) {
  r[0] = b[0] * c[0] - b[1] * c[1] - b[2] * c[2] - b[3] * c[3];
  r[1] = b[0] * c[1] + b[1] * c[0] + b[2] * c[3] - b[3] * c[2];
  r[2] = b[0] * c[2] + b[2] * c[0] + b[3] * c[1] - b[1] * c[3];
  r[3] = b[0] * c[3] + b[3] * c[0] + b[1] * c[2] - b[2] * c[1];
}

// ---------------------------------------------------------------- RNG
__device__ __forceinline__ void philox(uint64_t seed, uint32_t env, uint32_t episode, uint32_t tick, uint32_t stream, uint32_t* out) {
  uint32_t c0 = env, c1 = episode, c2 = tick, c3 = stream, k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int i = 0; i < 10; i++) {
    uint32_t h0 = __umulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
    uint32_t h1 = __umulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
    uint32_t n0 = h1 ^ c1 ^ k0, n2 = h0 ^ c3 ^ k1;
    c0 = n0; c1 = l1; c2 = n2; c3 = l0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
__device__ __forceinline__ float u11(uint32_t x) { return 2.0f * ((float)(x >> 8) * (1.0f / 16777216.0f)) - 1.0f; }

// ODE's dRand LCG advanced k steps at once: seed' = A_k*seed + C_k (mod 2^32)
__device__ __forceinline__ uint32_t lcg_jump(uint32_t seed, uint32_t k) {
  uint32_t accA = 1u, accC = 0u, curA = 1664525u, curC = 1013904223u;
  while (k) {
    if (k & 1u) { accC = curA * accC + curC; accA = curA * accA; }
    curC = (curA + 1u) * curC; curA = curA * curA; k >>= 1;
  }
  return accA * seed + accC;
}
struct FyMagic { uint32_t v[MAXROWS + 4]; };
constexpr FyMagic make_fy_magic() { FyMagic t{}; for (int i = 1; i < MAXROWS + 4; i++) t.v[i] = 0xffffffffu / (uint32_t)i; return t; }
__constant__ FyMagic c_fy_magic = make_fy_magic();   // floor((2^32 - 1) / un), read with a warp-uniform index

// dRandInt with the modulo done by a reciprocal multiply: magic = floor((2^32 - 1) / un) makes umulhi(r, magic) fall short of
// floor(r / un) by at most 2, which two conditional subtractions repair -- exact for every r, and the generic 32-bit
// division (the longest dependent stretch of the Fisher-Yates draw) leaves the chain; magic depends on un only.
__device__ __forceinline__ int lcg_randint_magic(uint32_t& seed, uint32_t un, uint32_t magic) {
  seed = 1664525u * seed + 1013904223u;
  uint32_t r = seed;
  if (un <= 0x00010000u) { r ^= (r >> 16);
    if (un <= 0x00000100u) { r ^= (r >> 8);
      if (un <= 0x00000010u) { r ^= (r >> 4);
        if (un <= 0x00000004u) { r ^= (r >> 2);
          if (un <= 0x00000002u) r ^= (r >> 1); } } } }
  uint32_t rem = r - __umulhi(r, magic) * un;
  if (rem >= un) rem -= un;
  if (rem >= un) rem -= un;
  return (int)rem;
}
__device__ __forceinline__ int lcg_randint(uint32_t& seed, uint32_t un) {  // dRandInt
  seed = 1664525u * seed + 1013904223u;
  uint32_t r = seed;
  if (un <= 0x00010000u) { r ^= (r >> 16);
    if (un <= 0x00000100u) { r ^= (r >> 8);
      if (un <= 0x00000010u) { r ^= (r >> 4);
        if (un <= 0x00000004u) { r ^= (r >> 2);
          if (un <= 0x00000002u) r ^= (r >> 1); } } } }
  return (int)(r % un);
}

// ---------------------------------------------------------------- colliders
struct CG { float p[3], n[3], depth; };

__device__ __forceinline__ int col_sphere_plane(const float* c, float r, CG* out) {  // plane z = 0, normal +z
  float depth = 0.0f - c[2] + r;
  if (depth >= 0) { out->n[0] = 0; out->n[1] = 0; out->n[2] = 1; out->p[0] = c[0]; out->p[1] = c[1]; out->p[2] = c[2] - r; out->depth = depth; return 1; }
  return 0;
}
__device__ __forceinline__ int col_sphere_box(const float* c, float r, const float* bpos, const float* bR, const float* side, CG* out) {
  float p[3] = {c[0] - bpos[0], c[1] - bpos[1], c[2] - bpos[2]}, l[3], t[3];
  bool onborder = false;
  mulRtv(t, bR, p);
#pragma unroll
  for (int i = 0; i < 3; i++) {
    l[i] = side[i] * 0.5f;
    if (t[i] < -l[i]) { t[i] = -l[i]; onborder = true; }
    if (t[i] > l[i]) { t[i] = l[i]; onborder = true; }
  }
  if (!onborder) {
    float md = l[0] - fabsf(t[0]); int mini = 0;
#pragma unroll
    for (int i = 1; i < 3; i++) { float fd = l[i] - fabsf(t[i]); if (fd < md) { md = fd; mini = i; } }
    out->p[0] = c[0]; out->p[1] = c[1]; out->p[2] = c[2];
    float sg = (t[mini] > 0) ? 1.0f : -1.0f;
    out->n[0] = bR[0 + mini] * sg; out->n[1] = bR[3 + mini] * sg; out->n[2] = bR[6 + mini] * sg;
    out->depth = md + r; return 1;
  }
  float q[3], rr[3];
  mulRv(q, bR, t);
  rr[0] = p[0] - q[0]; rr[1] = p[1] - q[1]; rr[2] = p[2] - q[2];
  float d2 = dot3(rr, rr);
  float depth = r - sqrtf(d2);
  if (depth < 0) return 0;
  out->p[0] = q[0] + bpos[0]; out->p[1] = q[1] + bpos[1]; out->p[2] = q[2] + bpos[2];
  if (d2 > 0) { float il = 1.0f / sqrtf(d2); out->n[0] = rr[0] * il; out->n[1] = rr[1] * il; out->n[2] = rr[2] * il; }
  else { out->n[0] = 1; out->n[1] = 0; out->n[2] = 0; }
  out->depth = depth; return 1;
}
__device__ int col_cyl_plane(const float* cpos, const float* R, float radius, float length, CG* out) {  // plane z=0
  const float tol = 0.0001f;
  int cnt = 0;
  float v[3] = {R[2], R[5], R[8]};
  float s = length * 0.5f;
  float g1[3], g2[3];
#pragma unroll
  for (int i = 0; i < 3; i++) { g2[i] = v[i] * s + cpos[i]; g1[i] = v[i] * -s + cpos[i]; }
  s = v[2];
  if (s < 0) s += 1.0f; else s -= 1.0f;
  if (s < tol && s > -tol) {
    float P[3];
    s = 0.0f - g1[2];
    float t = 0.0f - g2[2];
    if (s >= t) { if (s >= 0) { P[0] = g1[0]; P[1] = g1[1]; P[2] = g1[2]; } else return 0; }
    else { if (t >= 0) { P[0] = g2[0]; P[1] = g2[1]; P[2] = g2[2]; } else return 0; }
    float V1[3], V2[3];
    if (v[0] < tol && v[0] > -tol) { V1[0] = v[0] + 1.0f; V1[1] = v[1]; V1[2] = v[2]; }
    else { V1[0] = v[0]; V1[1] = v[1] + 1.0f; V1[2] = v[2]; }
    cross3(V2, V1, v);
    t = sqrtf(dot3(V2, V2)); t = radius / t;
    V2[0] *= t; V2[1] *= t; V2[2] *= t;
    cross3(V1, V2, v);
    for (int k = 0; k < 4; k++) {
      const float* V = (k < 2) ? V1 : V2; float sg = (k & 1) ? -1.0f : 1.0f;
      float p[3] = {P[0] + sg * V[0], P[1] + sg * V[1], P[2] + sg * V[2]};
      float dep = 0.0f - p[2];
      if (dep > 0) { out[cnt].p[0] = p[0]; out[cnt].p[1] = p[1]; out[cnt].p[2] = p[2]; out[cnt].n[0] = 0; out[cnt].n[1] = 0; out[cnt].n[2] = 1; out[cnt].depth = dep; cnt++; }
    }
  } else {
    float t = v[2];
    float C[3] = {v[0] * t - 0.0f, v[1] * t - 0.0f, v[2] * t - 1.0f};
    s = sqrtf(dot3(C, C)); s = radius / s;
    C[0] *= s; C[1] *= s; C[2] *= s;
#pragma unroll
    for (int d = 0; d < 2; d++) {
      const float* g = d == 0 ? g1 : g2;
      float p[3] = {C[0] + g[0], C[1] + g[1], C[2] + g[2]};
      float dep = 0.0f - p[2];
      if (dep >= 0) { out[cnt].p[0] = p[0]; out[cnt].p[1] = p[1]; out[cnt].p[2] = p[2]; out[cnt].n[0] = 0; out[cnt].n[1] = 0; out[cnt].n[2] = 1; out[cnt].depth = dep; cnt++; }
    }
  }
  return cnt;
}
__device__ int col_box_plane(const float* bpos, const float* R, const float* side, CG* out) {  // plane z=0, <=3 contacts
  float Q[3] = {R[6], R[7], R[8]};
  float A[3] = {side[0] * Q[0], side[1] * Q[1], side[2] * Q[2]};
  float B[3] = {fabsf(A[0]), fabsf(A[1]), fabsf(A[2])};
  float depth = 0.0f + 0.5f * (B[0] + B[1] + B[2]) - bpos[2];
  if (depth < 0) return 0;
  float p[3] = {bpos[0], bpos[1], bpos[2]};
#pragma unroll
  for (int i = 0; i < 3; i++) {
    float sg = (A[i] > 0) ? -1.0f : 1.0f;
    p[0] += sg * 0.5f * side[i] * R[0 + i]; p[1] += sg * 0.5f * side[i] * R[3 + i]; p[2] += sg * 0.5f * side[i] * R[6 + i];
  }
  out[0].p[0] = p[0]; out[0].p[1] = p[1]; out[0].p[2] = p[2]; out[0].n[0] = 0; out[0].n[1] = 0; out[0].n[2] = 1; out[0].depth = depth;
  int ret = 1;
  int first, second;
  if (B[0] < B[1]) {
    if (B[2] < B[0]) { first = 2; second = (B[0] < B[1]) ? 0 : 1; }
    else { first = 0; second = (B[1] < B[2]) ? 1 : 2; }
  } else {
    if (B[2] < B[1]) { first = 2; second = (B[0] < B[1]) ? 0 : 1; }
    else { first = 1; second = (B[0] < B[2]) ? 0 : 2; }
  }
  for (int c = 0; c < 2; c++) {
    int sd = c == 0 ? first : second;
    float Bs = sd == 0 ? B[0] : (sd == 1 ? B[1] : B[2]);
    float As = sd == 0 ? A[0] : (sd == 1 ? A[1] : A[2]);
    float ss = sd == 0 ? side[0] : (sd == 1 ? side[1] : side[2]);
    depth -= Bs;
    if (depth < 0) return ret;
    float sg = (As > 0) ? 1.0f : -1.0f;
    out[ret].p[0] = p[0] + sg * ss * R[0 + sd]; out[ret].p[1] = p[1] + sg * ss * R[3 + sd]; out[ret].p[2] = p[2] + sg * ss * R[6 + sd];
    out[ret].n[0] = 0; out[ret].n[1] = 0; out[ret].n[2] = 1; out[ret].depth = depth; ret++;
  }
  return ret;
}
// sample point strictly inside a terrain box -> pushed out through the nearest face
__device__ __forceinline__ bool point_in_box(const float* p, const float* bpos, const float* bR, const float* side, float* nrm, float* depth) {
  float d[3] = {p[0] - bpos[0], p[1] - bpos[1], p[2] - bpos[2]}, t[3];
  mulRtv(t, bR, d);
  float best = CUDART_INF_F; int mini = -1;
#pragma unroll
  for (int i = 0; i < 3; i++) {
    float fd = side[i] * 0.5f - fabsf(t[i]);
    if (fd <= 0) return false;
    if (fd < best) { best = fd; mini = i; }
  }
  float tm = mini == 0 ? t[0] : (mini == 1 ? t[1] : t[2]);
  float sg = (tm > 0) ? 1.0f : -1.0f;
  nrm[0] = bR[0 + mini] * sg; nrm[1] = bR[3 + mini] * sg; nrm[2] = bR[6 + mini] * sg;
  *depth = best; return true;
}

// ---- heightfield: HF-local frame x (columns), y up, z (rows); world = (x + ox, -(z - w/2) + py, y + pz)
__device__ __forceinline__ void hf_to_local(const HfParams& hf, const float* w, float* l) {
  l[0] = w[0] - hf.pos[0] + hf.length * 0.5f; l[1] = w[2] - hf.pos[2]; l[2] = -(w[1] - hf.pos[1]) + hf.width * 0.5f;
}
__device__ __forceinline__ float hfH(const HfParams& hf, int ix, int iz) { return __ldg(hf.h + iz * hf.ncol + ix); }
__device__ __forceinline__ bool point_vs_hf(const HfParams& hf, const float* pw, float* nrm_w, float* depth) {
  float p[3]; hf_to_local(hf, pw, p);
  if (p[0] < 0 || p[2] < 0 || p[0] > hf.length || p[2] > hf.width) return false;
  if (p[1] > hf.hmax) return false;
  int ix = (int)floorf(p[0] * hf.inv_sx), iz = (int)floorf(p[2] * hf.inv_sz);
  if (ix > hf.ncol - 2) ix = hf.ncol - 2;
  if (iz > hf.nrow - 2) iz = hf.nrow - 2;
  float x0 = ix * hf.sx, z0 = iz * hf.sz;
  float u = (p[0] - x0) * hf.inv_sx, v = (p[2] - z0) * hf.inv_sz;
  float hA = hfH(hf, ix, iz), hB = hfH(hf, ix + 1, iz), hC = hfH(hf, ix, iz + 1), hD = hfH(hf, ix + 1, iz + 1);
  float n[3], w[3];
  if (u + v <= 1) { n[0] = -(hB - hA) * hf.sz; n[1] = hf.sx * hf.sz; n[2] = -(hC - hA) * hf.sx; w[0] = p[0] - x0; w[1] = p[1] - hA; w[2] = p[2] - z0; }
  else { n[0] = -(hD - hC) * hf.sz; n[1] = hf.sx * hf.sz; n[2] = -(hD - hB) * hf.sx; w[0] = p[0] - (x0 + hf.sx); w[1] = p[1] - hD; w[2] = p[2] - (z0 + hf.sz); }
  float il = 1.0f / sqrtf(dot3(n, n)); n[0] *= il; n[1] *= il; n[2] *= il;
  float dist = dot3(n, w);
  if (!(dist < 0)) return false;
  nrm_w[0] = n[0]; nrm_w[1] = -n[2]; nrm_w[2] = n[1];
  *depth = -dist; return true;
}

// sphere vs heightfield: local minima of the centre-to-surface distance (faces, convex edges, convex
// vertices); scan order and tests: DESIGN.md D-HF.  The scan is warp-cooperative: a lane first decides
// whether its sphere can reach the surface at all (hf_sphere_range, block maxima), then the whole warp
// works on one candidate sphere at a time -- lane = owner cell for the prism reject, lane = (surviving
// cell, feature) for the six feature tests -- and emits the first four contacts in the canonical order
// (cells iz-major; per owner vertex, x-edge, z-edge, diagonal, face0, face1).
__device__ __forceinline__ bool hf_sphere_range(const HfParams& hf, const float* cw, float r, float* c, int* rng) {
  hf_to_local(hf, cw, c);
  int ix0 = (int)floorf((c[0] - r) * hf.inv_sx), ix1 = (int)floorf((c[0] + r) * hf.inv_sx);
  int iz0 = (int)floorf((c[2] - r) * hf.inv_sz), iz1 = (int)floorf((c[2] + r) * hf.inv_sz);
  if (ix0 < 0) ix0 = 0;
  if (iz0 < 0) iz0 = 0;
  if (ix1 > hf.ncol - 1) ix1 = hf.ncol - 1;
  if (iz1 > hf.nrow - 1) iz1 = hf.nrow - 1;
  rng[0] = ix0; rng[1] = ix1; rng[2] = iz0; rng[3] = iz1;
  if (ix0 > ix1 || iz0 > iz1) return false;
  const int bx1 = min(ix1 + 1, hf.ncol - 1) >> 3, bz1 = min(iz1 + 1, hf.nrow - 1) >> 3;
  float zm = -1e30f;
  for (int bz = iz0 >> 3; bz <= bz1; bz++) for (int bx = ix0 >> 3; bx <= bx1; bx++) zm = fmaxf(zm, __ldg(hf.blockmax + bz * hf.nbx + bx));
  return !(c[1] - r > zm);
}

// one feature (f = 0 vertex, 1 x-edge, 2 z-edge, 3 diagonal, 4 face0, 5 face1) of owner cell (ix, iz) against
// the sphere (centre c in the HF frame): unit normal nl (HF frame) and centre distance along it
__device__ __forceinline__ bool hf_feature(const HfParams& hf, const float* c, float r, int ix, int iz, int f, float* nl, float* dist_out) {
  const bool hx = ix + 1 <= hf.ncol - 1, hz = iz + 1 <= hf.nrow - 1;
  const float r2 = r * r;
  const float x0 = ix * hf.sx, z0 = iz * hf.sz;
  const float hA = hfH(hf, ix, iz);
  const float hB = hx ? hfH(hf, ix + 1, iz) : hA, hC = hz ? hfH(hf, ix, iz + 1) : hA, hD = (hx && hz) ? hfH(hf, ix + 1, iz + 1) : hA;
  const float A[3] = {x0, hA, z0}, B[3] = {x0 + hf.sx, hB, z0}, C[3] = {x0, hC, z0 + hf.sz}, D[3] = {x0 + hf.sx, hD, z0 + hf.sz};
  if (f == 0) {
    float d[3] = {c[0] - A[0], c[1] - A[1], c[2] - A[2]};
    float d2 = dot3(d, d);
    if (!(d2 < r2 && d2 > 0 && d[1] > 0)) return false;
    bool ok = true;
    const int nx[6] = {1, -1, 0, 0, 1, -1}, nz[6] = {0, 0, 1, -1, -1, 1};
#pragma unroll
    for (int k = 0; k < 6; k++) {
      int jx = ix + nx[k], jz = iz + nz[k];
      if (jx < 0 || jz < 0 || jx > hf.ncol - 1 || jz > hf.nrow - 1) continue;
      float u[3] = {nx[k] * hf.sx, hfH(hf, jx, jz) - hA, nz[k] * hf.sz};
      if (dot3(d, u) > 0) ok = false;
    }
    if (!ok) return false;
    float l = sqrtf(d2);
    nl[0] = d[0] / l; nl[1] = d[1] / l; nl[2] = d[2] / l; *dist_out = l;
    return true;
  }
  if (f <= 3) {
    const float *P0, *P1, *Q1; float Q2[3]; bool has1, has2;
    if (f == 1) { if (!hx) return false; P0 = A; P1 = B; Q1 = C; has1 = hz; has2 = iz - 1 >= 0; Q2[0] = x0 + hf.sx; Q2[1] = has2 ? hfH(hf, ix + 1, iz - 1) : 0.0f; Q2[2] = z0 - hf.sz; }
    else if (f == 2) { if (!hz) return false; P0 = A; P1 = C; Q1 = B; has1 = hx; has2 = ix - 1 >= 0; Q2[0] = x0 - hf.sx; Q2[1] = has2 ? hfH(hf, ix - 1, iz + 1) : 0.0f; Q2[2] = z0 + hf.sz; }
    else { if (!(hx && hz)) return false; P0 = B; P1 = C; Q1 = A; has1 = true; has2 = true; Q2[0] = D[0]; Q2[1] = D[1]; Q2[2] = D[2]; }
    float ev[3] = {P1[0] - P0[0], P1[1] - P0[1], P1[2] - P0[2]};
    float w[3] = {c[0] - P0[0], c[1] - P0[1], c[2] - P0[2]};
    float ee = dot3(ev, ev);
    float t = dot3(w, ev) / ee;
    if (!(t > 0 && t < 1)) return false;
    float d[3] = {w[0] - t * ev[0], w[1] - t * ev[1], w[2] - t * ev[2]};
    float d2 = dot3(d, d);
    if (!(d2 < r2 && d2 > 0 && d[1] > 0)) return false;
    bool ok = true;
#pragma unroll
    for (int k = 0; k < 2; k++) {
      if (!(k == 0 ? has1 : has2)) continue;
      const float* Q = (k == 0) ? Q1 : Q2;
      float g[3] = {Q[0] - P0[0], Q[1] - P0[1], Q[2] - P0[2]};
      float s = dot3(g, ev) / ee;
      float gp[3] = {g[0] - s * ev[0], g[1] - s * ev[1], g[2] - s * ev[2]};
      if (dot3(d, gp) > 0) ok = false;
    }
    if (!ok) return false;
    float l = sqrtf(d2);
    nl[0] = d[0] / l; nl[1] = d[1] / l; nl[2] = d[2] / l; *dist_out = l;
    return true;
  }
  if (!(hx && hz)) return false;
  float n[3];
  if (f == 4) { n[0] = -(hB - hA) * hf.sz; n[1] = hf.sx * hf.sz; n[2] = -(hC - hA) * hf.sx; }
  else { n[0] = -(hD - hC) * hf.sz; n[1] = hf.sx * hf.sz; n[2] = -(hD - hB) * hf.sx; }
  float il = 1.0f / sqrtf(dot3(n, n)); n[0] *= il; n[1] *= il; n[2] *= il;
  const float* P0 = (f == 4) ? A : D;
  float w[3] = {c[0] - P0[0], c[1] - P0[1], c[2] - P0[2]};
  float dist = dot3(n, w);
  if (!(dist < r)) return false;
  float fx = c[0] - n[0] * dist, fz = c[2] - n[2] * dist;
  float u = (fx - x0) * hf.inv_sx, v = (fz - z0) * hf.inv_sz;
  bool inside = (f == 4) ? (u > 0 && v > 0 && u + v < 1) : (u < 1 && v < 1 && u + v > 1);
  if (!inside) return false;
  nl[0] = n[0]; nl[1] = n[1]; nl[2] = n[2]; *dist_out = dist;
  return true;
}

// Warp-cooperative scan of one sphere (all arguments warp-uniform).  Contacts go to cont[slot0 ...] (capacity maxc);
// returns the number generated (<= 4).
__device__ int col_sphere_hf_warp(const HfParams& hf, const float* cw, const float* c, float r, const int* rng, int lane,
                                  float (*cont)[8], int slot0, int maxc, int info) {
  const int ix0 = rng[0], ix1 = rng[1], iz0 = rng[2], iz1 = rng[3];
  const int ncx = ix1 - ix0 + 1, ncell = ncx * (iz1 - iz0 + 1);
  const float r2 = r * r;
  const unsigned lt = (1u << lane) - 1u;
  int cnt = 0;
  for (int base = 0; base < ncell; base += 32) {
    const int cell = base + lane;
    bool live = false;
    if (cell < ncell) {  // conservative reject against the prism that holds every feature this vertex owns
      const int iz = iz0 + cell / ncx, ix = ix0 + cell % ncx;
      const bool hx = ix + 1 <= hf.ncol - 1, hz = iz + 1 <= hf.nrow - 1;
      const float x0 = ix * hf.sx, z0 = iz * hf.sz;
      const float hA = hfH(hf, ix, iz);
      const float hB = hx ? hfH(hf, ix + 1, iz) : hA, hC = hz ? hfH(hf, ix, iz + 1) : hA, hD = (hx && hz) ? hfH(hf, ix + 1, iz + 1) : hA;
      float lo = fminf(fminf(hA, hB), fminf(hC, hD)), hi = fmaxf(fmaxf(hA, hB), fmaxf(hC, hD));
      float dx = fmaxf(fmaxf(x0 - c[0], c[0] - (x0 + (hx ? hf.sx : 0.0f))), 0.0f);
      float dz = fmaxf(fmaxf(z0 - c[2], c[2] - (z0 + (hz ? hf.sz : 0.0f))), 0.0f);
      float dy = fmaxf(fmaxf(lo - c[1], c[1] - hi), 0.0f);
      live = !(dx * dx + dy * dy + dz * dz >= r2);
    }
    const unsigned lm = __ballot_sync(FULLMASK, live);
    const int nitems = __popc(lm) * 6;
    for (int i0 = 0; i0 < nitems; i0 += 32) {
      const int item = i0 + lane;
      bool hit = false; float nl[3] = {0, 0, 0}, dist = 0;
      if (item < nitems) {
        const int k = item / 6, f = item - k * 6;
        const int cl = base + (int)__fns(lm, 0, k + 1);
        hit = hf_feature(hf, c, r, ix0 + cl % ncx, iz0 + cl / ncx, f, nl, &dist);
      }
      const unsigned hm = __ballot_sync(FULLMASK, hit);
      if (hit) {
        const int slot = cnt + __popc(hm & lt);
        if (slot < 4 && slot0 + slot < maxc) {
          float* d = cont[slot0 + slot];
          const float nw0 = nl[0], nw1 = -nl[2], nw2 = nl[1];
          d[0] = cw[0] - nw0 * r; d[1] = cw[1] - nw1 * r; d[2] = cw[2] - nw2 * r; d[3] = nw0; d[4] = nw1; d[5] = nw2;
          d[6] = r - dist; d[7] = __int_as_float(info);
        }
      }
      cnt += __popc(hm);
      if (cnt >= 4) return 4;
    }
  }
  return cnt;
}

// terrain kinds inside the kernel
enum { T_PLANE = 0, T_BOX = 1, T_HF = 2 };

// point-sample test against a terrain box / the heightfield
__device__ __forceinline__ bool sample_test(int tkind, const float* q, const float* bpos, const float* bR, const float* side, const HfParams& hf, float* nrm, float* dep) {
  if (tkind == T_BOX) return point_in_box(q, bpos, bR, side, nrm, dep);
  return point_vs_hf(hf, q, nrm, dep);
}

__device__ int col_cyl_samples(int tkind, const float* p, const float* R, float radius, float length, const float* bpos, const float* bR, const float* side, const HfParams& hf, CG* out) {
  const float cs[8][2] = {{1.f, 0.f}, {0.70710678118654752f, 0.70710678118654752f}, {0.f, 1.f}, {-0.70710678118654752f, 0.70710678118654752f},
                          {-1.f, 0.f}, {-0.70710678118654752f, -0.70710678118654752f}, {0.f, -1.f}, {0.70710678118654752f, -0.70710678118654752f}};
  int cnt = 0;
#pragma unroll 1
  for (int cap = 0; cap < 2; cap++) {
    float s = (cap == 0 ? -0.5f : 0.5f) * length;
    float best = 0; bool any = false; CG bc;
#pragma unroll 1
    for (int k = 0; k < 8; k++) {
      float lx = radius * cs[k][0], ly = radius * cs[k][1];
      float q[3];
#pragma unroll
      for (int i = 0; i < 3; i++) q[i] = p[i] + R[i * 3 + 0] * lx + R[i * 3 + 1] * ly + R[i * 3 + 2] * s;
      float nrm[3], dep;
      if (sample_test(tkind, q, bpos, bR, side, hf, nrm, &dep) && (!any || dep > best)) {
        any = true; best = dep; bc.p[0] = q[0]; bc.p[1] = q[1]; bc.p[2] = q[2]; bc.n[0] = nrm[0]; bc.n[1] = nrm[1]; bc.n[2] = nrm[2]; bc.depth = dep;
      }
    }
    if (any) out[cnt++] = bc;
  }
  return cnt;
}
__device__ int col_box_samples(int tkind, const float* p, const float* R, const float* dim, const float* bpos, const float* bR, const float* side, const HfParams& hf, CG* out) {
  int cnt = 0;
#pragma unroll 1
  for (int k = 0; k < 8 && cnt < 4; k++) {
    float hx = ((k & 1) ? 0.5f : -0.5f) * dim[0], hy = ((k & 2) ? 0.5f : -0.5f) * dim[1], hz = ((k & 4) ? 0.5f : -0.5f) * dim[2];
    float q[3];
#pragma unroll
    for (int i = 0; i < 3; i++) q[i] = p[i] + R[i * 3 + 0] * hx + R[i * 3 + 1] * hy + R[i * 3 + 2] * hz;
    float nrm[3], dep;
    if (sample_test(tkind, q, bpos, bR, side, hf, nrm, &dep)) {
      out[cnt].p[0] = q[0]; out[cnt].p[1] = q[1]; out[cnt].p[2] = q[2]; out[cnt].n[0] = nrm[0]; out[cnt].n[1] = nrm[1]; out[cnt].n[2] = nrm[2]; out[cnt].depth = dep; cnt++;
    }
  }
  return cnt;
}

// hinge angle (ODE getHingeAngle): bodies' quaternions from the shared table
__device__ __forceinline__ float hinge_angle(const WarpSm& S, int j) {
  const Model& M = g_model;
  int b1 = M.jb1[j], b2 = M.jb2[j];
  float qq[4], qerr[4];
  qmul_inv_a(qq, S.bq[b1], S.bq[b2]);
  qmul_inv_b(qerr, qq, M.qrel[j]);
  float cost2 = qerr[0];
  float sint2 = sqrtf(qerr[1] * qerr[1] + qerr[2] * qerr[2] + qerr[3] * qerr[3]);
  float d = dot3(qerr + 1, M.axis1[j]);
  float theta = (d >= 0) ? (2 * tfn::atan2(sint2, cost2)) : (2 * tfn::atan2(sint2, -cost2));
  if (theta > 3.14159265358979323846f) theta -= 6.28318530717958647692f;
  return -theta;
}
__device__ __forceinline__ float hinge_rate(const WarpSm& S, int j) {
  const Model& M = g_model;
  int b1 = M.jb1[j], b2 = M.jb2[j];
  float ax[3]; mulRv(ax, S.bR[b1], M.axis1[j]);
  return dot3(ax, &S.bvel[b1][3]) - dot3(ax, &S.bvel[b2][3]);
}

// base island of a body: {0,1,2,3}->0 {4,6,7}->1 {5,8,9}->2 {10}->3 {11}->4
__device__ __forceinline__ int body_group(int b) { return (int)((0x432211210000ULL >> (b * 4)) & 15ULL); }

}  // namespace tfode
