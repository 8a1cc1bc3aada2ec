// =============================================================================
// CPU ORACLE -- TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the
// product path; only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may build, load or call it.
//
// PARITY UNPINNED for this file: the arithmetic it restates lives in libode
// (Open Dynamics Engine, Debian `libode-dev` 0.14 / 0.13.1, double precision;
// reference pin: CMakeLists.txt:10, Dockerfile:1,6, ode/environment.hh:22),
// whose source and binary are absent from /root/reference and from this box, and
// the reference has no tests or golden vectors.  What follows is a restatement of
// ODE's published algorithm for exactly the API slice the reference calls
// (SURVEY.md Appendix B), anchored on the reference's call sites:
//   dWorldQuickStep         <- ode/environment.hh:132-133
//   dCollide                <- ode/environment.cc:104
//   dJointCreateContact     <- ode/environment.cc:110-128
//   hinge / slider          <- ode/servo.cc:12-25, src/rover_1.cc:128-132,194-200,224-229, ode/object.cc:155-179
//   dMassSet*Total          <- ode/box.hh:57, ode/wheel.hh:60
//   dRFromEulerAngles       <- ode/object.cc:119
// Colliders cylinder-box, box-box and everything against the heightfield are
// specified afresh (DESIGN.md "Deviations"); the plane colliders, sphere-box,
// the joints and QuickStep follow ODE's algorithms.
// =============================================================================
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>

#include "../include/tfode_numerics.h"
#include "../include/tfode_lcp.h"

namespace ode_restated {

template <class T> inline T inf() { return std::numeric_limits<T>::infinity(); }

// ---------------------------------------------------------------- vec helpers
// The fma placement of dot/cross/matrix-vector products is part of the numerics contract
// (include/tfode_numerics.h); everything else below is plain IEEE arithmetic in source order.
template <class T> inline T dot3(const T* a, const T* b) { return tfn::dot3(a, b); }
template <class T> inline void cross3(T* r, const T* a, const T* b) { tfn::cross3(r, a, b); }
template <class T> inline void set3(T* r, T x, T y, T z) { r[0] = x; r[1] = y; r[2] = z; }
template <class T> inline void copy3(T* r, const T* a) { r[0] = a[0]; r[1] = a[1]; r[2] = a[2]; }
// R is 3x3 row-major (ODE pads rows to 4; the padding carries no arithmetic).
template <class T> inline void mulRv(T* r, const T* R, const T* v) { tfn::mulRv(r, R, v, 3); }   // dMultiply0_331
template <class T> inline void mulRtv(T* r, const T* R, const T* v) { tfn::mulRtv(r, R, v, 3); } // dMultiply1_331
// symmetric 3x3 (xx xy xz yy yz zz) times vector
template <class T> inline void symMulV(T* r, const T* S, const T* v) {
  T a = tfn::dot3(S[0], S[1], S[2], v[0], v[1], v[2]);
  T b = tfn::dot3(S[1], S[3], S[4], v[0], v[1], v[2]);
  T c = tfn::dot3(S[2], S[4], S[5], v[0], v[1], v[2]);
  r[0] = a; r[1] = b; r[2] = c;
}
template <class T> inline bool normalize3(T* a) {
  T l = dot3(a, a);
  if (l > 0) { l = T(1) / std::sqrt(l); a[0] *= l; a[1] *= l; a[2] *= l; return true; }
  a[0] = 1; a[1] = 0; a[2] = 0; return false;
}
template <class T> inline void normalize4(T* q) {  // dNormalize4
  T l = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
  if (l > 0) { l = T(1) / std::sqrt(l); q[0] *= l; q[1] *= l; q[2] *= l; q[3] *= l; }
  else { q[0] = 1; q[1] = q[2] = q[3] = 0; }
}

// ------------------------------------------------------------------ rotations
template <class T> inline void RFromEulerAngles(T* R, T phi, T theta, T psi) {  // ODE rotation.cpp
  T sphi = std::sin(phi), cphi = std::cos(phi), stheta = std::sin(theta), ctheta = std::cos(theta),
    spsi = std::sin(psi), cpsi = std::cos(psi);
  R[0] = cpsi * ctheta;  R[1] = spsi * ctheta;  R[2] = -stheta;
  R[3] = cpsi * stheta * sphi - spsi * cphi;  R[4] = spsi * stheta * sphi + cpsi * cphi;  R[5] = ctheta * sphi;
  R[6] = cpsi * stheta * cphi + spsi * sphi;  R[7] = spsi * stheta * cphi - cpsi * sphi;  R[8] = ctheta * cphi;
}
template <class T> inline void QfromR(T* q, const T* R) {  // dRtoQ
  T tr = R[0] + R[4] + R[8], s;
  if (tr >= 0) {
    s = std::sqrt(tr + 1); q[0] = T(0.5) * s; s = T(0.5) / s;
    q[1] = (R[7] - R[5]) * s; q[2] = (R[2] - R[6]) * s; q[3] = (R[3] - R[1]) * s;
  } else {
    if (R[4] > R[0]) {
      if (R[8] > R[4]) goto case_2;
      s = std::sqrt((R[4] - (R[8] + R[0])) + 1); q[2] = T(0.5) * s; s = T(0.5) / s;
      q[3] = (R[5] + R[7]) * s; q[1] = (R[3] + R[1]) * s; q[0] = (R[2] - R[6]) * s; return;
    }
    if (R[8] > R[0]) {
    case_2:
      s = std::sqrt((R[8] - (R[0] + R[4])) + 1); q[3] = T(0.5) * s; s = T(0.5) / s;
      q[1] = (R[6] + R[2]) * s; q[2] = (R[5] + R[7]) * s; q[0] = (R[3] - R[1]) * s; return;
    }
    s = std::sqrt((R[0] - (R[4] + R[8])) + 1); q[1] = T(0.5) * s; s = T(0.5) / s;
    q[2] = (R[3] + R[1]) * s; q[3] = (R[6] + R[2]) * s; q[0] = (R[7] - R[5]) * s;
  }
}
template <class T> inline void RfromQ(T* R, const T* q) {  // dQtoR
  T qq1 = 2 * q[1] * q[1], qq2 = 2 * q[2] * q[2], qq3 = 2 * q[3] * q[3];
  R[0] = 1 - qq2 - qq3;                R[1] = 2 * (q[1] * q[2] - q[0] * q[3]);  R[2] = 2 * (q[1] * q[3] + q[0] * q[2]);
  R[3] = 2 * (q[1] * q[2] + q[0] * q[3]);  R[4] = 1 - qq1 - qq3;                R[5] = 2 * (q[2] * q[3] - q[0] * q[1]);
  R[6] = 2 * (q[1] * q[3] - q[0] * q[2]);  R[7] = 2 * (q[2] * q[3] + q[0] * q[1]);  R[8] = 1 - qq1 - qq2;
}
template <class T> inline void QMul0(T* qa, const T* qb, const T* qc) {  // qa = qb*qc
  T a = qb[0] * qc[0] - qb[1] * qc[1] - qb[2] * qc[2] - qb[3] * qc[3];
  T b = qb[0] * qc[1] + qb[1] * qc[0] + qb[2] * qc[3] - qb[3] * qc[2];
  T c = qb[0] * qc[2] + qb[2] * qc[0] + qb[3] * qc[1] - qb[1] * qc[3];
  T d = qb[0] * qc[3] + qb[3] * qc[0] + qb[1] * qc[2] - qb[2] * qc[1];
  qa[0] = a; qa[1] = b; qa[2] = c; qa[3] = d;
}
template <class T> inline void QMul1(T* qa, const T* qb, const T* qc) {  // qa = qb^-1 * qc
  T a = qb[0] * qc[0] + qb[1] * qc[1] + qb[2] * qc[2] + qb[3] * qc[3];
  T b = qb[0] * qc[1] - qb[1] * qc[0] - qb[2] * qc[3] + qb[3] * qc[2];
  T c = qb[0] * qc[2] - qb[2] * qc[0] - qb[3] * qc[1] + qb[1] * qc[3];
  T d = qb[0] * qc[3] - qb[3] * qc[0] - qb[1] * qc[2] + qb[2] * qc[1];
  qa[0] = a; qa[1] = b; qa[2] = c; qa[3] = d;
}
template <class T> inline void QMul2(T* qa, const T* qb, const T* qc) {  // qa = qb * qc^-1
  T a = qb[0] * qc[0] + qb[1] * qc[1] + qb[2] * qc[2] + qb[3] * qc[3];
  T b = -qb[0] * qc[1] + qb[1] * qc[0] - qb[2] * qc[3] + qb[3] * qc[2];
  T c = -qb[0] * qc[2] + qb[2] * qc[0] - qb[3] * qc[1] + qb[1] * qc[3];
  T d = -qb[0] * qc[3] + qb[3] * qc[0] - qb[1] * qc[2] + qb[2] * qc[1];
  qa[0] = a; qa[1] = b; qa[2] = c; qa[3] = d;
}
template <class T> inline void QMul3(T* qa, const T* qb, const T* qc) {  // qa = qb^-1 * qc^-1
  T a = qb[0] * qc[0] - qb[1] * qc[1] - qb[2] * qc[2] - qb[3] * qc[3];
  T b = -qb[0] * qc[1] - qb[1] * qc[0] + qb[2] * qc[3] - qb[3] * qc[2];
  T c = -qb[0] * qc[2] - qb[2] * qc[0] + qb[3] * qc[1] - qb[1] * qc[3];
  T d = -qb[0] * qc[3] - qb[3] * qc[0] + qb[1] * qc[2] - qb[2] * qc[1];
  qa[0] = a; qa[1] = b; qa[2] = c; qa[3] = d;
}
// dPlaneSpace
template <class T> inline void planeSpace(const T* n, T* p, T* q) {
  if (std::fabs(n[2]) > T(0.70710678118654752440)) {
    T a = n[1] * n[1] + n[2] * n[2];
    T k = T(1) / std::sqrt(a);
    p[0] = 0; p[1] = -n[2] * k; p[2] = n[1] * k;
    q[0] = a * k; q[1] = -n[0] * p[2]; q[2] = n[0] * p[1];
  } else {
    T a = n[0] * n[0] + n[1] * n[1];
    T k = T(1) / std::sqrt(a);
    p[0] = -n[1] * k; p[1] = n[0] * k; p[2] = 0;
    q[0] = -n[2] * p[1]; q[1] = n[2] * p[0]; q[2] = a * k;
  }
}

// ------------------------------------------------------------------- dRandInt
// ODE misc.cpp: seed = 1664525*seed + 1013904223 (mod 2^32); xor-fold + modulus.
inline uint32_t dRand(uint32_t& seed) { seed = 1664525u * seed + 1013904223u; return seed; }
inline int dRandInt(uint32_t& seed, int n) {
  const uint32_t un = (uint32_t)n;
  uint32_t r = dRand(seed);
  if (un <= 0x00010000u) {
    r ^= (r >> 16);
    if (un <= 0x00000100u) {
      r ^= (r >> 8);
      if (un <= 0x00000010u) {
        r ^= (r >> 4);
        if (un <= 0x00000004u) {
          r ^= (r >> 2);
          if (un <= 0x00000002u) r ^= (r >> 1);
        }
      }
    }
  }
  return (int)(r % un);
}

// ---------------------------------------------------------------------- bodies
template <class T> struct Body {
  T pos[3] = {0, 0, 0}, q[4] = {1, 0, 0, 0}, R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  T lvel[3] = {0, 0, 0}, avel[3] = {0, 0, 0}, facc[3] = {0, 0, 0}, tacc[3] = {0, 0, 0};
  T mass = 1, invMass = 1, I[3] = {1, 1, 1}, invI[3] = {1, 1, 1};  // body-frame inertia is diagonal for every mass the rover uses
  void setRotation(const T* Rin) {  // dBodySetRotation
    QfromR(q, Rin); normalize4(q); RfromQ(R, q);
  }
  void setMassBoxTotal(T m, T lx, T ly, T lz) {  // dMassSetBoxTotal
    mass = m; invMass = T(1) / m;
    I[0] = m / T(12) * (ly * ly + lz * lz); I[1] = m / T(12) * (lx * lx + lz * lz); I[2] = m / T(12) * (lx * lx + ly * ly);
    for (int i = 0; i < 3; i++) invI[i] = T(1) / I[i];
  }
  void setMassCylinderTotalZ(T m, T r, T l) {  // dMassSetCylinderTotal(direction=3)
    mass = m; invMass = T(1) / m;
    T r2 = r * r;
    T Iperp = m * (T(0.25) * r2 + (T(1) / T(12)) * l * l);
    I[0] = Iperp; I[1] = Iperp; I[2] = m * T(0.5) * r2;
    for (int i = 0; i < 3; i++) invI[i] = T(1) / I[i];
  }
  void relPointPos(const T* p, T* out) const { mulRv(out, R, p); out[0] += pos[0]; out[1] += pos[1]; out[2] += pos[2]; }  // dBodyGetRelPointPos
  void relPointVel(const T* p, T* out) const {  // dBodyGetRelPointVel
    T pr[3]; mulRv(pr, R, p); T c[3]; cross3(c, avel, pr);
    out[0] = lvel[0] + c[0]; out[1] = lvel[1] + c[1]; out[2] = lvel[2] + c[2];
  }
  void posRelPoint(const T* p, T* out) const { T d[3] = {p[0] - pos[0], p[1] - pos[1], p[2] - pos[2]}; mulRtv(out, R, d); }  // dBodyGetPosRelPoint
  void vectorToWorld(const T* v, T* out) const { mulRv(out, R, v); }
  void addTorque(const T* t) { tacc[0] += t[0]; tacc[1] += t[1]; tacc[2] += t[2]; }
  void addForceAtRelPos(const T* f, const T* prel) {  // dBodyAddForceAtRelPos
    T p[3]; mulRv(p, R, prel);
    facc[0] += f[0]; facc[1] += f[1]; facc[2] += f[2];
    T c[3]; cross3(c, p, f); tacc[0] += c[0]; tacc[1] += c[1]; tacc[2] += c[2];
  }
};

// ---------------------------------------------------------------------- joints
template <class T> struct LimitMotor {  // dxJointLimitMotor
  T vel = 0, fmax = 0, lostop = -inf<T>(), histop = inf<T>(), fudge_factor = 1, normal_cfm = 0, stop_erp = 0, stop_cfm = 0;
  int limit = 0; T limit_err = 0;
  void setFMax(T v) { if (v >= 0) fmax = v; }  // negative values are ignored by ODE
  int testRotationalLimit(T angle) {
    if (angle <= lostop) { limit = 1; limit_err = angle - lostop; return 1; }
    else if (angle >= histop) { limit = 2; limit_err = angle - histop; return 1; }
    else { limit = 0; return 0; }
  }
};

enum JointType { J_HINGE = 0, J_SLIDER = 1, J_CONTACT = 2 };

template <class T> struct Joint {
  int type = J_HINGE;
  int b1 = -1, b2 = -1;  // body indices, -1 = world
  bool reverse = false;  // dJOINT_REVERSE
  T anchor1[3] = {0, 0, 0}, anchor2[3] = {0, 0, 0};
  T axis1[3] = {1, 0, 0}, axis2[3] = {1, 0, 0};
  T qrel[4] = {1, 0, 0, 0};
  T offset[3] = {0, 0, 0};  // slider
  LimitMotor<T> limot;
  // contact
  T cpos[3] = {0, 0, 0}, cnormal[3] = {0, 0, 1}, cdepth = 0, mu = 0, soft_erp = 0, soft_cfm = 0;
  bool soft = false;
};

struct Row12 { };

template <class T> struct World {
  std::vector<Body<T>> body;
  std::vector<Joint<T>> joint;    // persistent joints, creation order
  std::vector<Joint<T>> contact;  // contact joint group, emptied after every step
  T gravity[3] = {0, 0, 0};
  T erp = T(0.2), cfm = T(1e-10);
  int iters = 20; T sor_w = T(1.3);
  uint32_t rand_seed = 0;
  int solver = 0;           // 0: dWorldQuickStep (ode/environment.hh:133), 1: dWorldStep's dense Dantzig LCP (:132), include/tfode_lcp.h
  int lcp_failures = 0;     // islands whose Dantzig drive stalled (lcp.cpp's "s <= 0" exit)
  int last_rows = 0, last_islands = 0;

  // --- joint construction (joint.cpp / hinge.cpp / slider.cpp) ---------------
  int createHinge(int b1, int b2) {
    Joint<T> j; j.type = J_HINGE; attach(j, b1, b2);
    j.limot.normal_cfm = cfm; j.limot.stop_erp = erp; j.limot.stop_cfm = cfm;
    joint.push_back(j); return (int)joint.size() - 1;
  }
  int createSlider(int b1, int b2) {
    Joint<T> j; j.type = J_SLIDER; attach(j, b1, b2);
    j.limot.normal_cfm = cfm; j.limot.stop_erp = erp; j.limot.stop_cfm = cfm;
    joint.push_back(j); return (int)joint.size() - 1;
  }
  static void attach(Joint<T>& j, int b1, int b2) {  // dJointAttach: body1 must be non-null
    if (b1 < 0 && b2 >= 0) { j.b1 = b2; j.b2 = -1; j.reverse = true; }
    else { j.b1 = b1; j.b2 = b2; j.reverse = false; }
  }
  void setHingeAnchor(int ji, T x, T y, T z) {  // setAnchors
    Joint<T>& j = joint[ji]; T p[3] = {x, y, z};
    body[j.b1].posRelPoint(p, j.anchor1);
    if (j.b2 >= 0) body[j.b2].posRelPoint(p, j.anchor2); else copy3(j.anchor2, p);
    computeInitialRelativeRotation(j);
  }
  void setAxes(Joint<T>& j, T x, T y, T z) {
    T a[3] = {x, y, z}; normalize3(a);
    mulRtv(j.axis1, body[j.b1].R, a);
    if (j.b2 >= 0) mulRtv(j.axis2, body[j.b2].R, a); else copy3(j.axis2, a);
  }
  void setHingeAxis(int ji, T x, T y, T z) { setAxes(joint[ji], x, y, z); computeInitialRelativeRotation(joint[ji]); }
  void computeInitialRelativeRotation(Joint<T>& j) {
    if (j.b2 >= 0) QMul1(j.qrel, body[j.b1].q, body[j.b2].q);
    else { j.qrel[0] = body[j.b1].q[0]; for (int i = 1; i < 4; i++) j.qrel[i] = -body[j.b1].q[i]; }
  }
  void setSliderAxis(int ji, T x, T y, T z) {
    Joint<T>& j = joint[ji];
    setAxes(j, x, y, z);
    // computeOffset
    if (j.b2 >= 0) {
      T c[3] = {body[j.b1].pos[0] - body[j.b2].pos[0], body[j.b1].pos[1] - body[j.b2].pos[1], body[j.b1].pos[2] - body[j.b2].pos[2]};
      mulRtv(j.offset, body[j.b2].R, c);
    } else copy3(j.offset, body[j.b1].pos);
    computeInitialRelativeRotation(j);
  }
  // --- hinge getters ---------------------------------------------------------
  T hingeAngle(const Joint<T>& j) const {  // getHingeAngle + getHingeAngleFromRelativeQuat
    T qerr[4];
    if (j.b2 >= 0) { T qq[4]; QMul1(qq, body[j.b1].q, body[j.b2].q); QMul2(qerr, qq, j.qrel); }
    else QMul3(qerr, body[j.b1].q, j.qrel);
    T cost2 = qerr[0];
    T sint2 = std::sqrt(qerr[1] * qerr[1] + qerr[2] * qerr[2] + qerr[3] * qerr[3]);
    T theta = (dot3(qerr + 1, j.axis1) >= 0) ? (2 * tfn::atan2(sint2, cost2)) : (2 * tfn::atan2(sint2, -cost2));
    if (theta > T(M_PI)) theta -= T(2 * M_PI);
    theta = -theta;
    return theta;
  }
  T hingeAngleRate(const Joint<T>& j) const {
    T axis[3]; mulRv(axis, body[j.b1].R, j.axis1);
    T rate = dot3(axis, body[j.b1].avel);
    if (j.b2 >= 0) rate -= dot3(axis, body[j.b2].avel);
    if (j.reverse) rate = -rate;
    return rate;
  }
  void addHingeTorque(const Joint<T>& j, T torque) {
    if (j.reverse) torque = -torque;
    T axis[3]; mulRv(axis, body[j.b1].R, j.axis1);
    axis[0] *= torque; axis[1] *= torque; axis[2] *= torque;
    body[j.b1].addTorque(axis);
    if (j.b2 >= 0) { T n[3] = {-axis[0], -axis[1], -axis[2]}; body[j.b2].addTorque(n); }
  }
  T sliderPosition(const Joint<T>& j) const {
    T ax1[3], q[3]; mulRv(ax1, body[j.b1].R, j.axis1);
    if (j.b2 >= 0) {
      mulRv(q, body[j.b2].R, j.offset);
      for (int i = 0; i < 3; i++) q[i] = body[j.b1].pos[i] - q[i] - body[j.b2].pos[i];
    } else {
      for (int i = 0; i < 3; i++) q[i] = body[j.b1].pos[i] - j.offset[i];
      if (j.reverse) { ax1[0] = -ax1[0]; ax1[1] = -ax1[1]; ax1[2] = -ax1[2]; }
    }
    return dot3(ax1, q);
  }

  // --- constraint rows --------------------------------------------------------
  struct Rows {
    std::vector<T> J, c, cfm, lo, hi; std::vector<int> findex, jb;
    int m = 0;
    int add(int n, int b1, int b2, T world_cfm) {
      int r0 = m; m += n;
      J.resize((size_t)m * 12, T(0)); c.resize(m, T(0)); cfm.resize(m, world_cfm);
      lo.resize(m, -inf<T>()); hi.resize(m, inf<T>()); findex.resize(m, -1); jb.resize((size_t)m * 2);
      for (int i = r0; i < m; i++) { jb[i * 2] = b1; jb[i * 2 + 1] = b2; }
      return r0;
    }
    T* J1l(int r) { return &J[(size_t)r * 12]; }
    T* J1a(int r) { return &J[(size_t)r * 12 + 3]; }
    T* J2l(int r) { return &J[(size_t)r * 12 + 6]; }
    T* J2a(int r) { return &J[(size_t)r * 12 + 9]; }
  };

  // getInfo1: number of rows of a joint this step (also latches limot.limit)
  int info1(Joint<T>& j) {
    if (j.type == J_CONTACT) return j.mu > 0 ? 3 : 1;
    if (j.type == J_HINGE) {
      int m = (j.limot.fmax > 0) ? 6 : 5;
      j.limot.limit = 0;
      if ((j.limot.lostop >= -T(M_PI) || j.limot.histop <= T(M_PI)) && j.limot.lostop <= j.limot.histop) {
        T angle = hingeAngle(j);
        if (j.limot.testRotationalLimit(angle)) m = 6;
      }
      return m;
    }
    // slider
    int m = (j.limot.fmax > 0) ? 6 : 5;
    j.limot.limit = 0;
    if ((j.limot.lostop > -inf<T>() || j.limot.histop < inf<T>()) && j.limot.lostop <= j.limot.histop) {
      T pos = sliderPosition(j);
      if (j.limot.testRotationalLimit(pos)) m = 6;
    }
    return m;
  }

  // dxJointLimitMotor::addLimot
  void addLimot(Joint<T>& j, Rows& rw, int row, T fps, const T* ax1, bool rotational) {
    LimitMotor<T>& lm = j.limot;
    bool powered = lm.fmax > 0;
    if (!(powered || lm.limit)) return;
    T* J1 = rotational ? rw.J1a(row) : rw.J1l(row);
    T* J2 = rotational ? rw.J2a(row) : rw.J2l(row);
    copy3(J1, ax1);
    if (j.b2 >= 0) set3(J2, -ax1[0], -ax1[1], -ax1[2]);
    if (!rotational && j.b2 >= 0) {  // linear torque decoupling
      T c[3]; for (int i = 0; i < 3; i++) c[i] = T(0.5) * (body[j.b2].pos[i] - body[j.b1].pos[i]);
      T ltd[3]; cross3(ltd, c, ax1);
      copy3(rw.J1a(row), ltd); copy3(rw.J2a(row), ltd);
    }
    if (lm.limit && (lm.lostop == lm.histop)) powered = false;
    if (powered) {
      rw.cfm[row] = lm.normal_cfm;
      if (!lm.limit) { rw.c[row] = lm.vel; rw.lo[row] = -lm.fmax; rw.hi[row] = lm.fmax; }
      else {
        T fm = lm.fmax;
        if ((lm.vel > 0) || (lm.vel == 0 && lm.limit == 2)) fm = -fm;
        if ((lm.limit == 1 && lm.vel > 0) || (lm.limit == 2 && lm.vel < 0)) fm *= lm.fudge_factor;
        if (rotational) {
          T t[3] = {-fm * ax1[0], -fm * ax1[1], -fm * ax1[2]}; body[j.b1].addTorque(t);
          if (j.b2 >= 0) { T t2[3] = {fm * ax1[0], fm * ax1[1], fm * ax1[2]}; body[j.b2].addTorque(t2); }
        } else {
          for (int i = 0; i < 3; i++) body[j.b1].facc[i] += -fm * ax1[i];
          if (j.b2 >= 0) for (int i = 0; i < 3; i++) body[j.b2].facc[i] += fm * ax1[i];
        }
      }
    }
    if (lm.limit) {
      T k = fps * lm.stop_erp;
      rw.c[row] = -k * lm.limit_err;
      rw.cfm[row] = lm.stop_cfm;
      if (lm.lostop == lm.histop) { rw.lo[row] = -inf<T>(); rw.hi[row] = inf<T>(); }
      else if (lm.limit == 1) { rw.lo[row] = 0; rw.hi[row] = inf<T>(); }
      else { rw.lo[row] = -inf<T>(); rw.hi[row] = 0; }
    }
  }

  void setBall(Joint<T>& j, Rows& rw, int r0, T fps, T erp_) {
    T a1[3], a2[3];
    for (int i = 0; i < 3; i++) rw.J1l(r0 + i)[i] = 1;
    mulRv(a1, body[j.b1].R, j.anchor1);
    // J1a = -[a1]x
    rw.J1a(r0 + 0)[1] = a1[2];  rw.J1a(r0 + 0)[2] = -a1[1];
    rw.J1a(r0 + 1)[0] = -a1[2]; rw.J1a(r0 + 1)[2] = a1[0];
    rw.J1a(r0 + 2)[0] = a1[1];  rw.J1a(r0 + 2)[1] = -a1[0];
    T k = fps * erp_;
    if (j.b2 >= 0) {
      for (int i = 0; i < 3; i++) rw.J2l(r0 + i)[i] = -1;
      mulRv(a2, body[j.b2].R, j.anchor2);
      rw.J2a(r0 + 0)[1] = -a2[2]; rw.J2a(r0 + 0)[2] = a2[1];
      rw.J2a(r0 + 1)[0] = a2[2];  rw.J2a(r0 + 1)[2] = -a2[0];
      rw.J2a(r0 + 2)[0] = -a2[1]; rw.J2a(r0 + 2)[1] = a2[0];
      for (int i = 0; i < 3; i++) rw.c[r0 + i] = k * (a2[i] + body[j.b2].pos[i] - a1[i] - body[j.b1].pos[i]);
    } else {
      for (int i = 0; i < 3; i++) rw.c[r0 + i] = k * (j.anchor2[i] - a1[i] - body[j.b1].pos[i]);
    }
  }

  void setFixedOrientation(Joint<T>& j, Rows& rw, int r0, T fps, T erp_) {
    for (int i = 0; i < 3; i++) rw.J1a(r0 + i)[i] = 1;
    if (j.b2 >= 0) for (int i = 0; i < 3; i++) rw.J2a(r0 + i)[i] = -1;
    T qerr[4], e[3];
    if (j.b2 >= 0) { T qq[4]; QMul1(qq, body[j.b1].q, body[j.b2].q); QMul2(qerr, qq, j.qrel); }
    else QMul3(qerr, body[j.b1].q, j.qrel);
    if (qerr[0] < 0) { qerr[1] = -qerr[1]; qerr[2] = -qerr[2]; qerr[3] = -qerr[3]; }
    mulRv(e, body[j.b1].R, qerr + 1);
    T k = fps * erp_;
    for (int i = 0; i < 3; i++) rw.c[r0 + i] = 2 * k * e[i];
  }

  // getInfo2 for one joint; nrows from info1
  void info2(Joint<T>& j, Rows& rw, int r0, int nrows, T fps) {
    if (j.type == J_HINGE) {
      setBall(j, rw, r0, fps, erp);
      T ax1[3], p[3], q[3];
      mulRv(ax1, body[j.b1].R, j.axis1);
      planeSpace(ax1, p, q);
      copy3(rw.J1a(r0 + 3), p); copy3(rw.J1a(r0 + 4), q);
      if (j.b2 >= 0) { set3(rw.J2a(r0 + 3), -p[0], -p[1], -p[2]); set3(rw.J2a(r0 + 4), -q[0], -q[1], -q[2]); }
      T ax2[3], b[3];
      if (j.b2 >= 0) mulRv(ax2, body[j.b2].R, j.axis2); else copy3(ax2, j.axis2);
      cross3(b, ax1, ax2);
      T k = fps * erp;
      rw.c[r0 + 3] = k * dot3(b, p);
      rw.c[r0 + 4] = k * dot3(b, q);
      if (nrows == 6) addLimot(j, rw, r0 + 5, fps, ax1, true);
    } else if (j.type == J_SLIDER) {
      T c[3] = {0, 0, 0};
      const Body<T>& B1 = body[j.b1];
      if (j.b2 >= 0) for (int i = 0; i < 3; i++) c[i] = body[j.b2].pos[i] - B1.pos[i];
      setFixedOrientation(j, rw, r0, fps, erp);
      T ax1[3], p[3], q[3];
      mulRv(ax1, B1.R, j.axis1);
      planeSpace(ax1, p, q);
      if (j.b2 >= 0) {
        T tmp[3];
        cross3(tmp, c, p); for (int i = 0; i < 3; i++) { tmp[i] *= T(0.5); rw.J1a(r0 + 3)[i] = tmp[i]; rw.J2a(r0 + 3)[i] = tmp[i]; }
        cross3(tmp, c, q); for (int i = 0; i < 3; i++) { tmp[i] *= T(0.5); rw.J1a(r0 + 4)[i] = tmp[i]; rw.J2a(r0 + 4)[i] = tmp[i]; }
        for (int i = 0; i < 3; i++) { rw.J2l(r0 + 3)[i] = -p[i]; rw.J2l(r0 + 4)[i] = -q[i]; }
      }
      copy3(rw.J1l(r0 + 3), p); copy3(rw.J1l(r0 + 4), q);
      T k = fps * erp;
      if (j.b2 >= 0) {
        T ofs[3]; mulRv(ofs, body[j.b2].R, j.offset);
        for (int i = 0; i < 3; i++) c[i] += ofs[i];
        rw.c[r0 + 3] = k * dot3(p, c); rw.c[r0 + 4] = k * dot3(q, c);
      } else {
        T ofs[3]; for (int i = 0; i < 3; i++) ofs[i] = j.offset[i] - B1.pos[i];
        rw.c[r0 + 3] = k * dot3(p, ofs); rw.c[r0 + 4] = k * dot3(q, ofs);
        if (j.reverse) { ax1[0] = -ax1[0]; ax1[1] = -ax1[1]; ax1[2] = -ax1[2]; }
      }
      if (nrows == 6) addLimot(j, rw, r0 + 5, fps, ax1, false);
    } else {  // contact (dxJointContact::getInfo2, mode = dContactApprox1 [| SoftCFM | SoftERP])
      T normal[3];
      if (j.reverse) set3(normal, -j.cnormal[0], -j.cnormal[1], -j.cnormal[2]); else copy3(normal, j.cnormal);
      T c1[3], c2[3] = {0, 0, 0};
      for (int i = 0; i < 3; i++) c1[i] = j.cpos[i] - body[j.b1].pos[i];
      copy3(rw.J1l(r0), normal); cross3(rw.J1a(r0), c1, normal);
      if (j.b2 >= 0) {
        for (int i = 0; i < 3; i++) c2[i] = j.cpos[i] - body[j.b2].pos[i];
        set3(rw.J2l(r0), -normal[0], -normal[1], -normal[2]);
        T t[3]; cross3(t, c2, normal); set3(rw.J2a(r0), -t[0], -t[1], -t[2]);
      }
      T erp_ = j.soft ? j.soft_erp : erp;
      T k = fps * erp_;
      T depth = j.cdepth; if (depth < 0) depth = 0;  // contactp.min_depth = 0
      rw.c[r0] = k * depth;                           // contactp.max_vel = inf
      if (j.soft) rw.cfm[r0] = j.soft_cfm;
      rw.lo[r0] = 0; rw.hi[r0] = inf<T>();
      if (nrows >= 2) {
        T t1[3], t2[3]; planeSpace(normal, t1, t2);
        copy3(rw.J1l(r0 + 1), t1); cross3(rw.J1a(r0 + 1), c1, t1);
        if (j.b2 >= 0) { set3(rw.J2l(r0 + 1), -t1[0], -t1[1], -t1[2]); T t[3]; cross3(t, c2, t1); set3(rw.J2a(r0 + 1), -t[0], -t[1], -t[2]); }
        rw.lo[r0 + 1] = -j.mu; rw.hi[r0 + 1] = j.mu; rw.findex[r0 + 1] = r0;  // island-local index fixed up by the caller
        copy3(rw.J1l(r0 + 2), t2); cross3(rw.J1a(r0 + 2), c1, t2);
        if (j.b2 >= 0) { set3(rw.J2l(r0 + 2), -t2[0], -t2[1], -t2[2]); T t[3]; cross3(t, c2, t2); set3(rw.J2a(r0 + 2), -t[0], -t[1], -t[2]); }
        rw.lo[r0 + 2] = -j.mu; rw.hi[r0 + 2] = j.mu; rw.findex[r0 + 2] = r0;
      }
    }
  }

  // --- dWorldQuickStep ---------------------------------------------------------
  // Island order (our canonical definition, DESIGN.md "Deviations" D-ORDER): islands by
  // ascending smallest body index; inside an island joints in creation order, then
  // contacts in generation order.  ODE's own order depends on its intrusive list
  // layout and on the process-global dRand() state and cannot be reproduced.
  void quickstep(T h) {
    const int nb = (int)body.size();
    std::vector<int> parent(nb);
    for (int i = 0; i < nb; i++) parent[i] = i;
    auto find = [&](int x) { while (parent[x] != x) { parent[x] = parent[parent[x]]; x = parent[x]; } return x; };
    auto unite = [&](int a, int b) { a = find(a); b = find(b); if (a == b) return; if (a < b) parent[b] = a; else parent[a] = b; };
    for (auto& j : joint) if (j.b1 >= 0 && j.b2 >= 0) unite(j.b1, j.b2);
    for (auto& j : contact) if (j.b1 >= 0 && j.b2 >= 0) unite(j.b1, j.b2);
    last_rows = 0; last_islands = 0;
    for (int root = 0; root < nb; root++) {
      if (find(root) != root) continue;  // root = smallest index of its island (unite keeps the minimum)
      std::vector<int> ib; for (int i = 0; i < nb; i++) if (find(i) == root) ib.push_back(i);
      std::vector<Joint<T>*> ij;
      for (auto& j : joint) if (find(j.b1) == root) ij.push_back(&j);
      for (auto& j : contact) if (find(j.b1) == root) ij.push_back(&j);
      stepIsland(ib, ij, h);
      last_islands++;
    }
    for (auto& b : body) {  // dxStepBody + zero accumulators
      for (int k = 0; k < 3; k++) b.pos[k] += h * b.lvel[k];
      T dq[4];
      const T* w = b.avel; const T* q = b.q;
      dq[0] = T(0.5) * (-w[0] * q[1] - w[1] * q[2] - w[2] * q[3]);
      dq[1] = T(0.5) * (w[0] * q[0] + w[1] * q[3] - w[2] * q[2]);
      dq[2] = T(0.5) * (-w[0] * q[3] + w[1] * q[0] + w[2] * q[1]);
      dq[3] = T(0.5) * (w[0] * q[2] - w[1] * q[1] + w[2] * q[0]);
      for (int k = 0; k < 4; k++) b.q[k] += h * dq[k];
      normalize4(b.q); RfromQ(b.R, b.q);
      for (int k = 0; k < 3; k++) { b.facc[k] = 0; b.tacc[k] = 0; }
    }
    contact.clear();  // dJointGroupEmpty
  }

  // One island of dWorldQuickStep.  Arithmetic layout (the algebra is ODE's quickstep.cpp; the grouping of the
  // floating-point operations is this engine's numerics contract, shared with the CUDA kernel):
  //   invI_w  symmetric, 6 entries  S(r,c) = fma(R[r][2], i2*R[c][2], fma(R[r][1], i1*R[c][1], R[r][0]*(i0*R[c][0])))
  //   gyro    tau -= w x (R (Ib .* (R^T w)))
  //   rhs_i   = fma(c_i, 1/h, -(dot6(J1,t1) + dot6(J2,t2))),  t = f/m + v/h | invI tau + w/h
  //   Ad_i    = sor_w / (d1 + d2 + cfm_i/h),  d = fma(1/m, dot3(Jl,Jl), dot3(invI Ja, Ja))
  //   sweep   delta = fma(Ad, rhs - (dot6(J1,fc1) + dot6(J2,fc2)), -(lambda * (Ad*cfm/h)));  clamp;
  //           fc_lin = fma(delta/m, Jl, fc_lin);  fc_ang = fma(delta, invI Ja, fc_ang)
  // ODE pre-scales J and rhs by Ad and stores iMJ = M^-1 J^T; the form above is algebraically the same
  // update with J kept unscaled (half the per-row storage on the GPU).
  void stepIsland(const std::vector<int>& ib, const std::vector<Joint<T>*>& ij, T h) {
    const int nb = (int)ib.size();
    const T fps = T(1) / h;
    std::vector<int> local(body.size(), -1);
    for (int i = 0; i < nb; i++) local[ib[i]] = i;
    std::vector<T> invI((size_t)nb * 6);
    for (int i = 0; i < nb; i++) {
      Body<T>& b = body[ib[i]];
      T* S = &invI[(size_t)i * 6];
      const T* R = b.R;
      int n = 0;
      for (int r = 0; r < 3; r++) for (int c = r; c < 3; c++)
        S[n++] = tfn::fma(R[r * 3 + 2], b.invI[2] * R[c * 3 + 2], tfn::fma(R[r * 3 + 1], b.invI[1] * R[c * 3 + 1], R[r * 3 + 0] * (b.invI[0] * R[c * 3 + 0])));
      T wl[3]; mulRtv(wl, R, b.avel);
      wl[0] *= b.I[0]; wl[1] *= b.I[1]; wl[2] *= b.I[2];
      T Iw[3]; mulRv(Iw, R, wl);
      T cr[3]; cross3(cr, b.avel, Iw);
      for (int k = 0; k < 3; k++) b.tacc[k] -= cr[k];  // gyroscopic mode on (dBodyCreate default)
      for (int k = 0; k < 3; k++) b.facc[k] += b.mass * gravity[k];
    }
    // rows
    Rows rw;
    for (Joint<T>* jp : ij) {
      int n = info1(*jp);
      if (n <= 0) continue;
      int r0 = rw.add(n, jp->b1, jp->b2, cfm);
      info2(*jp, rw, r0, n, fps);
    }
    const int m = rw.m;
    last_rows += m;
    std::vector<T> cforce((size_t)nb * 6, T(0));
    if (m > 0) {
      for (int i = 0; i < m; i++) { rw.jb[i * 2] = local[rw.jb[i * 2]]; if (rw.jb[i * 2 + 1] >= 0) rw.jb[i * 2 + 1] = local[rw.jb[i * 2 + 1]]; }
      std::vector<T> tmp1((size_t)nb * 6), rhs(m);
      for (int i = 0; i < nb; i++) {
        Body<T>& b = body[ib[i]];
        for (int k = 0; k < 3; k++) tmp1[i * 6 + k] = b.facc[k] * b.invMass + b.lvel[k] * fps;
        T t[3]; symMulV(t, &invI[(size_t)i * 6], b.tacc);
        for (int k = 0; k < 3; k++) tmp1[i * 6 + 3 + k] = t[k] + b.avel[k] * fps;
      }
      for (int i = 0; i < m; i++) {
        const T* Jr = &rw.J[(size_t)i * 12];
        int b1 = rw.jb[i * 2], b2 = rw.jb[i * 2 + 1];
        T s = tfn::dot6(Jr, &tmp1[(size_t)b1 * 6]);
        if (b2 >= 0) s = s + tfn::dot6(Jr + 6, &tmp1[(size_t)b2 * 6]);
        rhs[i] = tfn::fma(rw.c[i], fps, -s);
        rw.cfm[i] *= fps;
      }
      std::vector<T> lambda(m, T(0));
      if (solver == 1) {   // dWorldStep: dense A = J M^-1 J^T + cfm/h, Dantzig LCP (shared with the CUDA kernel: include/tfode_lcp.h)
        std::vector<T> invM(nb), fw((size_t)tfl::island_fwork(m));
        std::vector<int> iw((size_t)tfl::lcp_iwork(m));
        for (int i = 0; i < nb; i++) invM[i] = body[ib[i]].invMass;
        lcp_failures += tfl::dantzig_island<T>(m, rw.J.data(), rw.jb.data(), rhs.data(), rw.cfm.data(), rw.lo.data(), rw.hi.data(), rw.findex.data(),
                                               invM.data(), invI.data(), lambda.data(), cforce.data(), fw.data(), iw.data());
      } else
      sorLcp(m, nb, rw, invI, ib, lambda, cforce, rhs);
    }
    // velocity update: first the constraint acceleration, then the external forces
    for (int i = 0; i < nb; i++) {
      Body<T>& b = body[ib[i]];
      for (int k = 0; k < 3; k++) { b.lvel[k] += h * cforce[i * 6 + k]; b.avel[k] += h * cforce[i * 6 + 3 + k]; }
      T hm = h * b.invMass;
      T ht[3];
      for (int k = 0; k < 3; k++) { b.lvel[k] += hm * b.facc[k]; ht[k] = b.tacc[k] * h; }
      T t[3]; symMulV(t, &invI[(size_t)i * 6], ht);
      for (int k = 0; k < 3; k++) b.avel[k] += t[k];
    }
  }

  // SOR_LCP of quickstep.cpp (RANDOMLY_REORDER_CONSTRAINTS, no warm starting)
  void sorLcp(int m, int nb, Rows& rw, const std::vector<T>& invI, const std::vector<int>& ib,
              std::vector<T>& lambda, std::vector<T>& fc, std::vector<T>& b) {
    (void)nb;
    std::vector<T> Ad(m), Adcfm(m);
    for (int i = 0; i < m; i++) {
      const T* Jr = &rw.J[(size_t)i * 12];
      int b1 = rw.jb[i * 2], b2 = rw.jb[i * 2 + 1];
      T a[3]; symMulV(a, &invI[(size_t)b1 * 6], Jr + 3);
      T diag = tfn::fma(body[ib[b1]].invMass, dot3(Jr, Jr), dot3(a, Jr + 3));
      if (b2 >= 0) {
        symMulV(a, &invI[(size_t)b2 * 6], Jr + 9);
        diag = diag + tfn::fma(body[ib[b2]].invMass, dot3(Jr + 6, Jr + 6), dot3(a, Jr + 9));
      }
      Ad[i] = sor_w / (diag + rw.cfm[i]);
      Adcfm[i] = Ad[i] * rw.cfm[i];
    }
    std::vector<int> order(m);
    { int j = 0, k = 1;
      for (int i = 0; i < m; i++) { if (rw.findex[i] < 0) order[j++] = i; else { order[m - k] = i; k++; } } }
    for (int it = 0; it < iters; it++) {
      if ((it & 7) == 0) {
        for (int i = 1; i < m; i++) { int s = dRandInt(rand_seed, i + 1); std::swap(order[i], order[s]); }
      }
      for (int i = 0; i < m; i++) {
        int idx = order[i];
        T lo_act, hi_act;
        if (rw.findex[idx] >= 0) { hi_act = std::fabs(rw.hi[idx] * lambda[rw.findex[idx]]); lo_act = -hi_act; }
        else { hi_act = rw.hi[idx]; lo_act = rw.lo[idx]; }
        int b1 = rw.jb[idx * 2], b2 = rw.jb[idx * 2 + 1];
        const T* Jr = &rw.J[(size_t)idx * 12];
        T old = lambda[idx];
        T* f1 = &fc[(size_t)b1 * 6];
        T* f2 = b2 >= 0 ? &fc[(size_t)b2 * 6] : nullptr;
        T sum = tfn::dot6(Jr, f1);
        if (f2) sum = sum + tfn::dot6(Jr + 6, f2);
        T delta = tfn::fma(Ad[idx], b[idx] - sum, -(old * Adcfm[idx]));
        T nl = old + delta;
        if (nl < lo_act) { delta = lo_act - old; lambda[idx] = lo_act; }
        else if (nl > hi_act) { delta = hi_act - old; lambda[idx] = hi_act; }
        else lambda[idx] = nl;
        {
          const T dm = delta * body[ib[b1]].invMass;
          T a[3]; symMulV(a, &invI[(size_t)b1 * 6], Jr + 3);
          for (int c = 0; c < 3; c++) { f1[c] = tfn::fma(dm, Jr[c], f1[c]); f1[3 + c] = tfn::fma(delta, a[c], f1[3 + c]); }
        }
        if (f2) {
          const T dm = delta * body[ib[b2]].invMass;
          T a[3]; symMulV(a, &invI[(size_t)b2 * 6], Jr + 9);
          for (int c = 0; c < 3; c++) { f2[c] = tfn::fma(dm, Jr[6 + c], f2[c]); f2[3 + c] = tfn::fma(delta, a[c], f2[3 + c]); }
        }
      }
    }
  }
};

// =============================================================================
// Colliders.  Contact convention: normal points from g2 (terrain) into g1 (rover
// geom); depth >= 0.  Each returns the number of contacts written (<= maxc).
// =============================================================================
template <class T> struct ContactGeom { T pos[3], normal[3], depth; };

// dCollideSpherePlane (plane: n.x = d)
template <class T> inline int collideSpherePlane(const T* c, T r, const T* n, T d, ContactGeom<T>* out) {
  T k = dot3(c, n);
  T depth = d - k + r;
  if (depth >= 0) {
    copy3(out->normal, n);
    for (int i = 0; i < 3; i++) out->pos[i] = c[i] - n[i] * r;
    out->depth = depth; return 1;
  }
  return 0;
}

// dCollideSphereBox
template <class T> inline int collideSphereBox(const T* c, T r, const T* bpos, const T* bR, const T* side, ContactGeom<T>* out) {
  T p[3] = {c[0] - bpos[0], c[1] - bpos[1], c[2] - bpos[2]}, l[3], t[3];
  bool onborder = false;
  mulRtv(t, bR, p);
  for (int i = 0; i < 3; i++) {
    l[i] = side[i] * T(0.5);
    if (t[i] < -l[i]) { t[i] = -l[i]; onborder = true; }
    if (t[i] > l[i]) { t[i] = l[i]; onborder = true; }
  }
  if (!onborder) {
    T min_distance = l[0] - std::fabs(t[0]); int mini = 0;
    for (int i = 1; i < 3; i++) { T fd = l[i] - std::fabs(t[i]); if (fd < min_distance) { min_distance = fd; mini = i; } }
    copy3(out->pos, c);
    T tmp[3] = {0, 0, 0}; tmp[mini] = (t[mini] > 0) ? T(1) : T(-1);
    mulRv(out->normal, bR, tmp);
    out->depth = min_distance + r; return 1;
  }
  T q[3], rr[3];
  mulRv(q, bR, t);
  for (int i = 0; i < 3; i++) rr[i] = p[i] - q[i];
  T depth = r - std::sqrt(dot3(rr, rr));
  if (depth < 0) return 0;
  for (int i = 0; i < 3; i++) out->pos[i] = q[i] + bpos[i];
  copy3(out->normal, rr); normalize3(out->normal);
  out->depth = depth; return 1;
}

// dCollideCylinderPlane (collision_cylinder_plane.cpp); axis = 3rd column of R
template <class T> inline int collideCylinderPlane(const T* cpos, const T* R, T radius, T length, const T* n, T d,
                                                   ContactGeom<T>* out, int maxc) {
  const T tol = (sizeof(T) == 4) ? T(0.0001) : T(0.0000001);
  int cnt = 0;
  T vDir1[3] = {R[2], R[5], R[8]};
  T s = length * T(0.5);
  T G1Pos1[3], G1Pos2[3];
  for (int i = 0; i < 3; i++) { G1Pos2[i] = vDir1[i] * s + cpos[i]; G1Pos1[i] = vDir1[i] * -s + cpos[i]; }
  s = dot3(vDir1, n);
  if (s < 0) s += T(1); else s -= T(1);
  auto emit = [&](const T* p, T depth) { copy3(out[cnt].pos, p); copy3(out[cnt].normal, n); out[cnt].depth = depth; cnt++; };
  if (s < tol && s > -tol) {
    T P[3];
    s = d - dot3(n, G1Pos1);
    T t = d - dot3(n, G1Pos2);
    if (s >= t) { if (s >= 0) copy3(P, G1Pos1); else return 0; }
    else { if (t >= 0) copy3(P, G1Pos2); else return 0; }
    T V1[3], V2[3];
    if (vDir1[0] < tol && vDir1[0] > -tol) { V1[0] = vDir1[0] + T(1); V1[1] = vDir1[1]; V1[2] = vDir1[2]; }
    else { V1[0] = vDir1[0]; V1[1] = vDir1[1] + T(1); V1[2] = vDir1[2]; }
    cross3(V2, V1, vDir1);
    t = std::sqrt(dot3(V2, V2)); t = radius / t;
    V2[0] *= t; V2[1] *= t; V2[2] *= t;
    cross3(V1, V2, vDir1);
    T p[3], dep;
    for (int k = 0; k < 4; k++) {
      const T* V = (k < 2) ? V1 : V2; T sg = (k & 1) ? T(-1) : T(1);
      for (int i = 0; i < 3; i++) p[i] = P[i] + sg * V[i];
      dep = d - dot3(n, p);
      if (dep > 0) { emit(p, dep); if (cnt >= maxc) return cnt; }
    }
  } else {
    T t = dot3(n, vDir1);
    T C[3] = {vDir1[0] * t - n[0], vDir1[1] * t - n[1], vDir1[2] * t - n[2]};
    s = std::sqrt(dot3(C, C)); s = radius / s;
    C[0] *= s; C[1] *= s; C[2] *= s;
    T p[3], dep;
    for (int i = 0; i < 3; i++) p[i] = C[i] + G1Pos1[i];
    dep = d - dot3(n, p);
    if (dep >= 0) { emit(p, dep); if (cnt >= maxc) return cnt; }
    for (int i = 0; i < 3; i++) p[i] = C[i] + G1Pos2[i];
    dep = d - dot3(n, p);
    if (dep >= 0) { emit(p, dep); if (cnt >= maxc) return cnt; }
  }
  return cnt;
}

// dCollideBoxPlane (box.cpp), at most 3 contacts
template <class T> inline int collideBoxPlane(const T* bpos, const T* R, const T* side, const T* n, T d, ContactGeom<T>* out, int maxc) {
  if (maxc > 3) maxc = 3;
  T Q[3] = {n[0] * R[0] + n[1] * R[3] + n[2] * R[6], n[0] * R[1] + n[1] * R[4] + n[2] * R[7], n[0] * R[2] + n[1] * R[5] + n[2] * R[8]};
  T A[3] = {side[0] * Q[0], side[1] * Q[1], side[2] * Q[2]};
  T B[3] = {std::fabs(A[0]), std::fabs(A[1]), std::fabs(A[2])};
  T depth = d + T(0.5) * (B[0] + B[1] + B[2]) - dot3(n, bpos);
  if (depth < 0) return 0;
  T p[3] = {bpos[0], bpos[1], bpos[2]};
  for (int i = 0; i < 3; i++) {
    T sg = (A[i] > 0) ? T(-1) : T(1);
    p[0] += sg * T(0.5) * side[i] * R[0 + i]; p[1] += sg * T(0.5) * side[i] * R[3 + i]; p[2] += sg * T(0.5) * side[i] * R[6 + i];
  }
  copy3(out[0].pos, p); copy3(out[0].normal, n); out[0].depth = depth;
  int ret = 1;
  if (maxc == 1) return ret;
  // second and third contacts: from p along the two sides with the smallest projected length
  int first, second;
  if (B[0] < B[1]) {
    if (B[2] < B[0]) { first = 2; second = (B[0] < B[1]) ? 0 : 1; }
    else { first = 0; second = (B[1] < B[2]) ? 1 : 2; }
  } else {
    if (B[2] < B[1]) { first = 2; second = (B[0] < B[1]) ? 0 : 1; }
    else { first = 1; second = (B[0] < B[2]) ? 0 : 2; }
  }
  int sides[2] = {first, second};
  for (int c = 0; c < 2 && ret < maxc; c++) {
    int sd = sides[c];
    depth -= B[sd];
    if (depth < 0) return ret;
    T sg = (A[sd] > 0) ? T(1) : T(-1);
    out[ret].pos[0] = p[0] + sg * side[sd] * R[0 + sd];
    out[ret].pos[1] = p[1] + sg * side[sd] * R[3 + sd];
    out[ret].pos[2] = p[2] + sg * side[sd] * R[6 + sd];
    copy3(out[ret].normal, n); out[ret].depth = depth; ret++;
  }
  return ret;
}

// ---- point vs box (specified afresh, DESIGN.md D-CYLBOX): a sample point of the rover
// geom that lies strictly inside the terrain box gives a contact at that point, pushed
// out through the nearest face.
template <class T> inline bool pointInBox(const T* p, const T* bpos, const T* bR, const T* side, T* normal, T* depth) {
  T d[3] = {p[0] - bpos[0], p[1] - bpos[1], p[2] - bpos[2]}, t[3];
  mulRtv(t, bR, d);
  T best = inf<T>(); int mini = -1;
  for (int i = 0; i < 3; i++) {
    T fd = side[i] * T(0.5) - std::fabs(t[i]);
    if (fd <= 0) return false;
    if (fd < best) { best = fd; mini = i; }
  }
  T tmp[3] = {0, 0, 0}; tmp[mini] = (t[mini] > 0) ? T(1) : T(-1);
  mulRv(normal, bR, tmp);
  *depth = best; return true;
}

// ---- heightfield (specified afresh, DESIGN.md D-HF) -------------------------
template <class T> struct HeightField {
  const float* h = nullptr;  // h[iz*ncol+ix], metres
  int ncol = 0, nrow = 0;
  T sx = 0, sz = 0, inv_sx = 0, inv_sz = 0, length = 0, width = 0;
  T pos[3] = {0, 0, 0};
  T hmin = 0, hmax = 0;
  std::vector<float> blockmax;  // 8x8-vertex block maxima (conservative early-out only)
  int nbx = 0, nbz = 0;
  void init(const float* data, int ncol_, int nrow_, double length_, double width_, const double* p) {
    h = data; ncol = ncol_; nrow = nrow_; length = (T)length_; width = (T)width_;
    sx = (T)(length_ / (ncol - 1)); sz = (T)(width_ / (nrow - 1));
    inv_sx = (T)((ncol - 1) / length_); inv_sz = (T)((nrow - 1) / width_);
    for (int i = 0; i < 3; i++) pos[i] = (T)p[i];
    nbx = (ncol + 7) / 8; nbz = (nrow + 7) / 8;
    blockmax.assign((size_t)nbx * nbz, -1e30f);
    float mn = 1e30f, mx = -1e30f;
    for (int iz = 0; iz < nrow; iz++) for (int ix = 0; ix < ncol; ix++) {
      float v = h[iz * ncol + ix]; mn = std::min(mn, v); mx = std::max(mx, v);
      float& b = blockmax[(size_t)(iz / 8) * nbx + ix / 8]; b = std::max(b, v);
    }
    hmin = mn; hmax = mx;
  }
  // world -> HF local: x along columns, z along rows (world -y), y up (world z)
  void toLocal(const T* w, T* l) const { l[0] = w[0] - pos[0] + length * T(0.5); l[1] = w[2] - pos[2]; l[2] = -(w[1] - pos[1]) + width * T(0.5); }
  void vecToWorld(const T* l, T* w) const { w[0] = l[0]; w[1] = -l[2]; w[2] = l[1]; }
  void pointToWorld(const T* l, T* w) const { w[0] = l[0] + pos[0] - length * T(0.5); w[1] = -(l[2] - width * T(0.5)) + pos[1]; w[2] = l[1] + pos[2]; }
  T H(int ix, int iz) const { return (T)h[iz * ncol + ix]; }
  float zoneMax(int ix0, int ix1, int iz0, int iz1) const {
    float m = -1e30f;
    for (int bz = iz0 / 8; bz <= iz1 / 8; bz++) for (int bx = ix0 / 8; bx <= ix1 / 8; bx++) m = std::max(m, blockmax[(size_t)bz * nbx + bx]);
    return m;
  }
};

// sphere vs heightfield: one contact per local minimum of the distance from the centre to
// the triangulated surface (faces, convex edges, convex vertices), at most maxc, scanned
// iz-major over the owner vertices; per owner: vertex, x-edge, z-edge, diagonal, face0, face1.
template <class T> inline int collideSphereHF(const T* cw, T r, const HeightField<T>& hf, ContactGeom<T>* out, int maxc) {
  T c[3]; hf.toLocal(cw, c);
  int ix0 = (int)std::floor((c[0] - r) * hf.inv_sx), ix1 = (int)std::floor((c[0] + r) * hf.inv_sx);
  int iz0 = (int)std::floor((c[2] - r) * hf.inv_sz), iz1 = (int)std::floor((c[2] + r) * hf.inv_sz);
  if (ix0 < 0) ix0 = 0; if (iz0 < 0) iz0 = 0;
  if (ix1 > hf.ncol - 1) ix1 = hf.ncol - 1; if (iz1 > hf.nrow - 1) iz1 = hf.nrow - 1;
  if (ix0 > ix1 || iz0 > iz1) return 0;
  if (c[1] - r > (T)hf.zoneMax(ix0, std::min(ix1 + 1, hf.ncol - 1), iz0, std::min(iz1 + 1, hf.nrow - 1))) return 0;
  const T r2 = r * r;
  int cnt = 0;
  auto emit = [&](const T* nl, T dist) {  // nl: unit normal in HF frame; dist: centre distance along it
    T nw[3]; hf.vecToWorld(nl, nw);
    copy3(out[cnt].normal, nw);
    for (int i = 0; i < 3; i++) out[cnt].pos[i] = cw[i] - nw[i] * r;
    out[cnt].depth = r - dist; cnt++;
  };
  for (int iz = iz0; iz <= iz1; iz++) for (int ix = ix0; ix <= ix1; ix++) {
    const bool hx = ix + 1 <= hf.ncol - 1, hz = iz + 1 <= hf.nrow - 1;
    const T x0 = ix * hf.sx, z0 = iz * hf.sz;
    const T hA = hf.H(ix, iz);
    const T hB = hx ? hf.H(ix + 1, iz) : hA, hC = hz ? hf.H(ix, iz + 1) : hA, hD = (hx && hz) ? hf.H(ix + 1, iz + 1) : hA;
    {  // conservative reject against the prism that holds every feature this vertex owns
      T lo = std::min(std::min(hA, hB), std::min(hC, hD)), hi = std::max(std::max(hA, hB), std::max(hC, hD));
      T dx = std::max(std::max(x0 - c[0], c[0] - (x0 + (hx ? hf.sx : T(0)))), T(0));
      T dz = std::max(std::max(z0 - c[2], c[2] - (z0 + (hz ? hf.sz : T(0)))), T(0));
      T dy = std::max(std::max(lo - c[1], c[1] - hi), T(0));
      if (dx * dx + dy * dy + dz * dz >= r2) continue;
    }
    const T A[3] = {x0, hA, z0}, B[3] = {x0 + hf.sx, hB, z0}, C[3] = {x0, hC, z0 + hf.sz}, D[3] = {x0 + hf.sx, hD, z0 + hf.sz};
    // -- vertex A
    {
      T d[3] = {c[0] - A[0], c[1] - A[1], c[2] - A[2]};
      T d2 = dot3(d, d);
      if (d2 < r2 && d2 > 0 && d[1] > 0) {
        bool ok = true;
        // six neighbours: (+1,0) (-1,0) (0,+1) (0,-1) (+1,-1) (-1,+1)
        const int nx[6] = {1, -1, 0, 0, 1, -1}, nz[6] = {0, 0, 1, -1, -1, 1};
        for (int k = 0; k < 6 && ok; k++) {
          int jx = ix + nx[k], jz = iz + nz[k];
          if (jx < 0 || jz < 0 || jx > hf.ncol - 1 || jz > hf.nrow - 1) continue;
          T u[3] = {nx[k] * hf.sx, hf.H(jx, jz) - hA, nz[k] * hf.sz};
          if (dot3(d, u) > 0) ok = false;
        }
        if (ok) { T l = std::sqrt(d2); T n[3] = {d[0] / l, d[1] / l, d[2] / l}; emit(n, l); if (cnt >= maxc) return cnt; }
      }
    }
    // -- edges: P0->P1 with the third vertices of the (up to two) adjacent faces
    auto edge = [&](const T* P0, const T* P1, bool has1, const T* Q1, bool has2, const T* Q2) -> bool {
      T e[3] = {P1[0] - P0[0], P1[1] - P0[1], P1[2] - P0[2]};
      T w[3] = {c[0] - P0[0], c[1] - P0[1], c[2] - P0[2]};
      T ee = dot3(e, e);
      T t = dot3(w, e) / ee;
      if (!(t > 0 && t < 1)) return false;
      T d[3] = {w[0] - t * e[0], w[1] - t * e[1], w[2] - t * e[2]};
      T d2 = dot3(d, d);
      if (!(d2 < r2 && d2 > 0 && d[1] > 0)) return false;
      for (int f = 0; f < 2; f++) {
        if (!(f == 0 ? has1 : has2)) continue;
        const T* Q = (f == 0) ? Q1 : Q2;
        T g[3] = {Q[0] - P0[0], Q[1] - P0[1], Q[2] - P0[2]};
        T s = dot3(g, e) / ee;
        T gp[3] = {g[0] - s * e[0], g[1] - s * e[1], g[2] - s * e[2]};
        if (dot3(d, gp) > 0) return false;
      }
      T l = std::sqrt(d2); T n[3] = {d[0] / l, d[1] / l, d[2] / l};
      emit(n, l); return true;
    };
    if (hx) {  // x-edge A->B: faces F0 of this cell (third C) and F1 of cell (ix, iz-1) (third vertex (ix+1, iz-1))
      bool f2 = iz - 1 >= 0;
      T Q2[3] = {x0 + hf.sx, f2 ? hf.H(ix + 1, iz - 1) : T(0), z0 - hf.sz};
      if (edge(A, B, hz, C, f2, Q2) && cnt >= maxc) return cnt;
    }
    if (hz) {  // z-edge A->C: faces F0 of this cell (third B) and F1 of cell (ix-1, iz) (third vertex (ix-1, iz+1))
      bool f2 = ix - 1 >= 0;
      T Q2[3] = {x0 - hf.sx, f2 ? hf.H(ix - 1, iz + 1) : T(0), z0 + hf.sz};
      if (edge(A, C, hx, B, f2, Q2) && cnt >= maxc) return cnt;
    }
    if (hx && hz) {
      if (edge(B, C, true, A, true, D) && cnt >= maxc) return cnt;  // diagonal
      // -- faces: foot of the perpendicular strictly inside the triangle
      for (int f = 0; f < 2; f++) {
        T n[3];
        if (f == 0) { n[0] = -(hB - hA) * hf.sz; n[1] = hf.sx * hf.sz; n[2] = -(hC - hA) * hf.sx; }
        else { n[0] = -(hD - hC) * hf.sz; n[1] = hf.sx * hf.sz; n[2] = -(hD - hB) * hf.sx; }
        normalize3(n);
        const T* P0 = (f == 0) ? A : D;
        T w[3] = {c[0] - P0[0], c[1] - P0[1], c[2] - P0[2]};
        T dist = dot3(n, w);
        if (!(dist < r)) continue;
        T fx = c[0] - n[0] * dist, fz = c[2] - n[2] * dist;
        T u = (fx - x0) * hf.inv_sx, v = (fz - z0) * hf.inv_sz;
        bool inside = (f == 0) ? (u > 0 && v > 0 && u + v < 1) : (u < 1 && v < 1 && u + v > 1);
        if (!inside) continue;
        emit(n, dist); if (cnt >= maxc) return cnt;
      }
    }
  }
  return cnt;
}

// a sample point of a rover geom against the triangulated surface: contact when the point is
// below the triangle it projects onto (depth along that triangle's normal)
template <class T> inline bool pointVsHF(const T* pw, const HeightField<T>& hf, T* normal_w, T* depth) {
  T p[3]; hf.toLocal(pw, p);
  if (p[0] < 0 || p[2] < 0 || p[0] > hf.length || p[2] > hf.width) return false;
  if (p[1] > hf.hmax) return false;
  int ix = (int)std::floor(p[0] * hf.inv_sx), iz = (int)std::floor(p[2] * hf.inv_sz);
  if (ix > hf.ncol - 2) ix = hf.ncol - 2; if (iz > hf.nrow - 2) iz = hf.nrow - 2;
  T x0 = ix * hf.sx, z0 = iz * hf.sz;
  T u = (p[0] - x0) * hf.inv_sx, v = (p[2] - z0) * hf.inv_sz;
  T hA = hf.H(ix, iz), hB = hf.H(ix + 1, iz), hC = hf.H(ix, iz + 1), hD = hf.H(ix + 1, iz + 1);
  T n[3], w[3];
  if (u + v <= 1) { n[0] = -(hB - hA) * hf.sz; n[1] = hf.sx * hf.sz; n[2] = -(hC - hA) * hf.sx; w[0] = p[0] - x0; w[1] = p[1] - hA; w[2] = p[2] - z0; }
  else { n[0] = -(hD - hC) * hf.sz; n[1] = hf.sx * hf.sz; n[2] = -(hD - hB) * hf.sx; w[0] = p[0] - (x0 + hf.sx); w[1] = p[1] - hD; w[2] = p[2] - (z0 + hf.sz); }
  normalize3(n);
  T dist = dot3(n, w);
  if (!(dist < 0)) return false;
  hf.vecToWorld(n, normal_w);
  *depth = -dist; return true;
}

}  // namespace ode_restated
