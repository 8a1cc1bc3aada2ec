/*
 * tfode_lcp.h -- the dense mixed-LCP island solver behind TFODE_SOLVER_DANTZIG: the equivalent of what
 * dWorldStep does for one island (ode/environment.hh:132, the call every shipped TF_ODE binary makes):
 *
 *     A = J M^-1 J^T + diag(cfm / h),   A lambda = rhs + w,   lo <= lambda <= hi,   w complementary to the bounds,
 *     friction rows coupled to their normal row through findex (bounds +-mu * lambda_normal, dContactApprox1),
 *
 * solved with Dantzig's pivoting method in the form ODE's lcp.cpp (dSolveLCP) uses: the unbounded rows are
 * factorised first, the remaining rows are driven to feasibility one at a time (non-friction rows before friction
 * rows), each drive moving the clamped set C / the at-bound set N by one index per pivot.
 *
 * One header for both sides, so that the CUDA kernel and the CPU oracle execute the same operations in the same
 * order: instantiated with float by tf_ode_b200/csrc (device code, -fmad=false) and with float / double by
 * oracle/ode_restated.hpp (-ffp-contract=off).  No fused multiply-adds are placed here; + - * / sqrt are correctly
 * rounded on both sides, so the float instantiations agree bit for bit.
 *
 * PARITY UNPINNED against libode (absent): the pivoting rules follow lcp.cpp as published, but the
 * factor of A_CC is a Cholesky factor kept in clamping order and rebuilt from the removed position when an index
 * leaves C (lcp.cpp updates an LDL^T factor with dLDLTAddTL / dLDLTRemove), and the row order is this engine's
 * canonical one (DESIGN.md Deviations, D-ORDER) -- mathematically the same pivots, not the same rounding.
 */
#ifndef TFODE_LCP_H_
#define TFODE_LCP_H_

#include "tfode_numerics.h"

namespace tfl {

template <class T> TFN_HD T t_inf();
template <> TFN_HD float t_inf<float>() { return __builtin_huge_valf(); }
template <> TFN_HD double t_inf<double>() { return __builtin_huge_val(); }
TFN_HD float t_sqrt(float x) { return sqrtf(x); }
TFN_HD double t_sqrt(double x) { return sqrt(x); }
template <class T> TFN_HD T t_abs(T x) { return x < T(0) ? -x : x; }
template <class T> TFN_HD T t_eps();
template <> TFN_HD float t_eps<float>() { return 1.1920929e-7f; }
template <> TFN_HD double t_eps<double>() { return 2.220446049250313e-16; }

/* work sizes for an island of m rows */
TFN_HD int lcp_fwork(int m) { return 2 * m * m + 8 * m; }
TFN_HD int lcp_iwork(int m) { return 3 * m; }

enum { ST_NEW = 0, ST_C = 1, ST_NLO = 2, ST_NHI = 3, ST_DEP = 4 /* dependent on the clamped rows: lambda stays 0 */ };

template <class T> struct Lcp {
  int n, nC;
  const T* A;   /* n x n, row-major */
  T* L;         /* Cholesky factor of A(C,C), rows/columns in clamping order, row-major n x n */
  T *x, *w, *lo, *hi, *dx, *dw, *ya;
  const T* b;
  int *Cl, *st, *pos;

  /* row `k` of L for index r = Cl[k] (rows 0..k-1 are valid) */
  TFN_HD void chol_row(int k) {
    const int r = Cl[k];
    T d = A[r * n + r];
    for (int j = 0; j < k; j++) {
      T s = A[r * n + Cl[j]];
      for (int t = 0; t < j; t++) s = s - L[k * n + t] * L[j * n + t];
      s = s / L[j * n + j];
      L[k * n + j] = s;
      d = d - s * s;
    }
    if (!(d > T(0))) d = T(1e-30);   /* A is positive definite (cfm > 0); this only guards rounding on degenerate input */
    L[k * n + k] = t_sqrt(d);
  }
  TFN_HD void to_C(int i) { Cl[nC] = i; pos[i] = nC; st[i] = ST_C; chol_row(nC); nC++; }
  TFN_HD void from_C(int i) {
    const int p = pos[i];
    for (int k = p; k + 1 < nC; k++) { Cl[k] = Cl[k + 1]; pos[Cl[k]] = k; }
    nC--;
    for (int k = p; k < nC; k++) chol_row(k);
  }
  /* ya = A(C,C)^-1 * v(C), v given as column `col` of A (col >= 0) or as the vector b (col < 0) */
  TFN_HD void solve_C(int col) {
    for (int j = 0; j < nC; j++) {
      T s = col >= 0 ? A[Cl[j] * n + col] : b[Cl[j]];
      for (int t = 0; t < j; t++) s = s - L[j * n + t] * ya[t];
      ya[j] = s / L[j * n + j];
    }
    for (int j = nC - 1; j >= 0; j--) {
      T s = ya[j];
      for (int t = j + 1; t < nC; t++) s = s - L[t * n + j] * ya[t];
      ya[j] = s / L[j * n + j];
    }
  }
  /* A(i,:) . x over the indexes that can be non-zero (C and N) */
  TFN_HD T row_dot_x(int i) const {
    T s = T(0);
    for (int k = 0; k < n; k++) if (st[k] != ST_NEW) s = s + A[i * n + k] * x[k];
    return s;
  }
};

/*
 * Solve the mixed LCP.  lo_in / hi_in: bounds; for a friction row (findex >= 0) hi_in holds mu and the bounds become
 * +-|mu * x[findex]| once every non-friction row is solved.  Returns 0, or 1 when a drive stalled (step <= 0 or the
 * pivot budget ran out): the remaining rows then keep lambda = 0, as lcp.cpp does ("cross our fingers and exit").
 * fwork: lcp_fwork(n) values, iwork: lcp_iwork(n).
 */
template <class T>
TFN_HD int lcp_dantzig(int n, const T* A, T* x, const T* b, const T* lo_in, const T* hi_in, const int* findex, T* fwork, int* iwork) {
  Lcp<T> P;
  P.n = n; P.nC = 0; P.A = A; P.b = b; P.x = x;
  P.L = fwork; P.w = fwork + n * n; P.lo = P.w + n; P.hi = P.lo + n; P.dx = P.hi + n; P.dw = P.dx + n; P.ya = P.dw + n;
  P.Cl = iwork; P.st = iwork + n; P.pos = iwork + 2 * n;
  const T inf = t_inf<T>();
  for (int i = 0; i < n; i++) { x[i] = T(0); P.w[i] = T(0); P.lo[i] = lo_in[i]; P.hi[i] = hi_in[i]; P.st[i] = ST_NEW; P.pos[i] = -1; }
  /* 1. every unbounded row goes to C first; x(C) = A(C,C)^-1 b(C) */
  for (int i = 0; i < n; i++) if (findex[i] < 0 && P.lo[i] == -inf && P.hi[i] == inf) P.to_C(i);
  if (P.nC > 0) { P.solve_C(-1); for (int j = 0; j < P.nC; j++) x[P.Cl[j]] = P.ya[j]; }
  int status = 0, budget = 8 * n + 16;
  /* 2. the bounded rows: pass 0 the non-friction ones in row order, pass 1 the friction ones */
  for (int pass = 0; pass < 2 && !status; pass++) {
    if (pass == 1) {
      for (int k = 0; k < n; k++) if (findex[k] >= 0) {
        const T wfk = x[findex[k]];
        if (wfk == T(0)) { P.hi[k] = T(0); P.lo[k] = T(0); } else { P.hi[k] = t_abs(P.hi[k] * wfk); P.lo[k] = -P.hi[k]; }
      }
    }
    for (int i = 0; i < n && !status; i++) {
      if (P.st[i] != ST_NEW || (findex[i] >= 0) != (pass == 1)) continue;
      P.w[i] = P.row_dot_x(i) - b[i];
      if (P.lo[i] == T(0) && P.w[i] >= T(0)) { P.st[i] = ST_NLO; continue; }
      if (P.hi[i] == T(0) && P.w[i] <= T(0)) { P.st[i] = ST_NHI; continue; }
      if (P.w[i] == T(0)) { P.to_C(i); continue; }
      for (;;) {   /* push x(i), w(i) to their equilibrium */
        const T dirf = P.w[i] <= T(0) ? T(1) : T(-1);
        P.solve_C(i);   /* delta_x(C) = -dir * A(C,C)^-1 A(C,i) */
        for (int j = 0; j < P.nC; j++) P.dx[P.Cl[j]] = -dirf * P.ya[j];
        for (int k = 0; k < n; k++) {
          if (k != i && P.st[k] != ST_NLO && P.st[k] != ST_NHI) continue;
          T s = T(0);
          for (int j = 0; j < P.nC; j++) s = s + A[k * n + P.Cl[j]] * P.dx[P.Cl[j]];
          P.dw[k] = s + A[k * n + i] * dirf;
        }
        /* A row that is a linear combination of the clamped rows (two tire spheres of one wheel give two lateral friction rows through
           almost the same point) has dw(i) = A(i,i) - A(i,C) A(C,C)^-1 A(C,i) = cfm/h in exact arithmetic -- 1e-7 for the reference's CFM --
           and rounding noise of either sign in binary32.  Such a row cannot be driven: it keeps lambda = 0 and the rows it depends on
           carry its share (J^T lambda, hence the motion, is the same for any split among dependent rows).  The threshold is 64 ulp of
           A(i,i): never reached in binary64, where cfm/h itself resolves the pivot as libode's double build does. */
        if (!(P.dw[i] * dirf > T(64) * t_eps<T>() * A[i * n + i])) { x[i] = T(0); P.st[i] = ST_DEP; break; }
        int cmd = 1, si = 0;
        T s = -P.w[i] / P.dw[i];
        if (dirf > T(0)) { if (P.hi[i] < inf) { const T s2 = (P.hi[i] - x[i]) * dirf; if (s2 < s) { s = s2; cmd = 3; } } }
        else { if (P.lo[i] > -inf) { const T s2 = (P.lo[i] - x[i]) * dirf; if (s2 < s) { s = s2; cmd = 2; } } }
        for (int k = 0; k < n; k++) {
          if (!((P.st[k] == ST_NLO && P.dw[k] < T(0)) || (P.st[k] == ST_NHI && P.dw[k] > T(0)))) continue;
          if (P.lo[k] == T(0) && P.hi[k] == T(0)) continue;
          const T s2 = -P.w[k] / P.dw[k];
          if (s2 < s) { s = s2; cmd = 4; si = k; }
        }
        for (int j = 0; j < P.nC; j++) {
          const int k = P.Cl[j];
          if (P.dx[k] < T(0) && P.lo[k] > -inf) { const T s2 = (P.lo[k] - x[k]) / P.dx[k]; if (s2 < s) { s = s2; cmd = 5; si = k; } }
          if (P.dx[k] > T(0) && P.hi[k] < inf) { const T s2 = (P.hi[k] - x[k]) / P.dx[k]; if (s2 < s) { s = s2; cmd = 6; si = k; } }
        }
#ifdef TFL_DEBUG
        if (!(s >= T(0)) || (s == T(0) && cmd < 4) || budget <= 0) fprintf(stderr, "LCP stall: n=%d i=%d findex=%d nC=%d cmd=%d si=%d s=%g w[i]=%g dw[i]=%g x[i]=%g lo=%g hi=%g dirf=%g budget=%d\n", n, i, findex[i], P.nC, cmd, si, (double)s, (double)P.w[i], (double)P.dw[i], (double)x[i], (double)P.lo[i], (double)P.hi[i], (double)dirf, budget);
#endif
        /* lcp.cpp gives up on s <= 0.  A zero-length pivot is a legitimate degenerate step though (an index that entered C sitting exactly
           on its bound and is asked to move outward goes straight back to N), so only a negative step -- or an exhausted pivot budget,
           which also bounds any cycling of such steps -- abandons the drive. */
        if (!(s >= T(0)) || (s == T(0) && cmd < 4) || --budget < 0) { x[i] = T(0); P.w[i] = T(0); status = 1; break; }   /* stalled: give up on the remaining rows, as lcp.cpp does */
        for (int j = 0; j < P.nC; j++) { const int k = P.Cl[j]; x[k] = x[k] + s * P.dx[k]; }
        x[i] = x[i] + s * dirf;
        for (int k = 0; k < n; k++) if (P.st[k] == ST_NLO || P.st[k] == ST_NHI) P.w[k] = P.w[k] + s * P.dw[k];
        P.w[i] = P.w[i] + s * P.dw[i];
        if (cmd == 1) { P.w[i] = T(0); P.to_C(i); break; }
        if (cmd == 2) { x[i] = P.lo[i]; P.st[i] = ST_NLO; break; }
        if (cmd == 3) { x[i] = P.hi[i]; P.st[i] = ST_NHI; break; }
        if (cmd == 4) { P.w[si] = T(0); P.to_C(si); }
        else if (cmd == 5) { x[si] = P.lo[si]; P.from_C(si); P.st[si] = ST_NLO; }
        else { x[si] = P.hi[si]; P.from_C(si); P.st[si] = ST_NHI; }
      }
    }
  }
  return status;
}

/*
 * One island of dWorldStep after the rows are built: A = J M^-1 J^T + diag(cfm) (cfm already divided by h), the LCP,
 * and the constraint acceleration fc = M^-1 J^T lambda accumulated per body.
 *   J[m][12] = J1l J1a J2l J2a;  jb[m][2] = body indexes into invM / invI / fc (second < 0: no body);
 *   invI[body][6] = xx xy xz yy yz zz (world frame);  fc[body][6] is ADDED to.
 * fwork: m*m + lcp_fwork(m) + 12*m values, iwork: lcp_iwork(m).
 */
TFN_HD int island_fwork(int m) { return m * m + lcp_fwork(m) + 12 * m; }
template <class T>
TFN_HD int dantzig_island(int m, const T* J, const int* jb, const T* rhs, const T* cfm, const T* lo, const T* hi, const int* findex,
                          const T* invM, const T* invI, T* lambda, T* fc, T* fwork, int* iwork) {
  T* A = fwork; T* iMJ = A + m * m; T* rest = iMJ + 12 * m;
  for (int i = 0; i < m; i++) {
    for (int sd = 0; sd < 2; sd++) {
      const int bd = jb[i * 2 + sd];
      T* o = iMJ + i * 12 + sd * 6; const T* Jr = J + i * 12 + sd * 6;
      if (bd < 0) { for (int c = 0; c < 6; c++) o[c] = T(0); continue; }
      const T* S = invI + bd * 6;
      o[0] = invM[bd] * Jr[0]; o[1] = invM[bd] * Jr[1]; o[2] = invM[bd] * Jr[2];
      o[3] = S[0] * Jr[3] + S[1] * Jr[4] + S[2] * Jr[5];
      o[4] = S[1] * Jr[3] + S[3] * Jr[4] + S[4] * Jr[5];
      o[5] = S[2] * Jr[3] + S[4] * Jr[4] + S[5] * Jr[5];
    }
  }
  for (int i = 0; i < m; i++) for (int j = i; j < m; j++) {
    T s = T(0);
    for (int si = 0; si < 2; si++) for (int sj = 0; sj < 2; sj++) {
      const int bi = jb[i * 2 + si];
      if (bi < 0 || bi != jb[j * 2 + sj]) continue;
      const T* a = iMJ + i * 12 + si * 6; const T* q = J + j * 12 + sj * 6;
      s = s + (a[0] * q[0] + a[1] * q[1] + a[2] * q[2] + a[3] * q[3] + a[4] * q[4] + a[5] * q[5]);
    }
    if (i == j) s = s + cfm[i];
    A[i * m + j] = s; A[j * m + i] = s;
  }
  const int status = lcp_dantzig(m, A, lambda, rhs, lo, hi, findex, rest, iwork);
  for (int i = 0; i < m; i++) for (int sd = 0; sd < 2; sd++) {
    const int bd = jb[i * 2 + sd];
    if (bd < 0) continue;
    const T* a = iMJ + i * 12 + sd * 6; T* f = fc + bd * 6;
    for (int c = 0; c < 6; c++) f[c] = f[c] + a[c] * lambda[i];
  }
  return status;
}

}  /* namespace tfl */
#endif /* TFODE_LCP_H_ */
