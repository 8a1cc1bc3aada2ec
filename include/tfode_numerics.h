/*
 * tfode_numerics.h -- the numerics contract of the batched rover stepping engine.
 *
 * The engine computes in IEEE-754 binary32 with round-to-nearest-even, no flush-to-zero,
 * correctly rounded + - * / sqrt and explicitly placed fused multiply-adds.  Both the CUDA
 * kernels (compiled with -fmad=false, so that nvcc never contracts on its own) and the CPU
 * oracle's float instantiation (compiled with -ffp-contract=off) write their arithmetic with
 * the primitives below, which makes a GPU trajectory reproducible bit for bit on a CPU --
 * the property the parity tests assert (free-running trajectories and contact counts
 * bit-exact against oracle<float>).
 *
 * Only primitives live here: fma-ordered dot/cross products and the handful of
 * transcendental functions the path needs (the reference calls libm / Eigen / libode for
 * them: atan2 in dJointGetHingeAngle and Eigen::AngleAxisd, ode/ft_sensor.cc:73; tan in
 * src/rover_1.cc:357-360; asin in src/rover_1.cc:262-299; tanh in the actor's output layer,
 * scripts/rover_training_1.py:33).  libm and the CUDA math library round these differently,
 * so the contract pins them to fixed polynomial evaluations (Cephes single-precision
 * kernels) built from the correctly rounded operations only.  T = double forwards to libm:
 * the fp64 oracle is the reference-equivalent arithmetic and carries no bit-exactness claim.
 */
#ifndef TFODE_NUMERICS_H_
#define TFODE_NUMERICS_H_

#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define TFN_HD __host__ __device__ __forceinline__
#else
#define TFN_HD inline
#endif

namespace tfn {

// ---------------------------------------------------------------- fused multiply-add, a*b + c with one rounding
TFN_HD float fma(float a, float b, float c) {
#if defined(__CUDA_ARCH__)
  return __fmaf_rn(a, b, c);
#else
  return ::fmaf(a, b, c);
#endif
}
TFN_HD double fma(double a, double b, double c) {
#if defined(__CUDA_ARCH__)
  return __fma_rn(a, b, c);
#else
  return ::fma(a, b, c);
#endif
}

// a0*b0 + a1*b1 + a2*b2 as  fma(a2,b2, fma(a1,b1, a0*b0))
template <class T> TFN_HD T dot3(const T* a, const T* b) { return fma(a[2], b[2], fma(a[1], b[1], a[0] * b[0])); }
template <class T> TFN_HD T dot3(T a0, T a1, T a2, T b0, T b1, T b2) { return fma(a2, b2, fma(a1, b1, a0 * b0)); }
// six-term dot product (a Jacobian half-row against a body's 6-vector): linear triple + angular triple
template <class T> TFN_HD T dot6(const T* a, const T* b) { return dot3(a, b) + dot3(a + 3, b + 3); }
// a x b with each component as fma(a_i, b_j, -(a_j * b_i))
template <class T> TFN_HD void cross3(T* r, const T* a, const T* b) {
  T x = fma(a[1], b[2], -(a[2] * b[1]));
  T y = fma(a[2], b[0], -(a[0] * b[2]));
  T z = fma(a[0], b[1], -(a[1] * b[0]));
  r[0] = x; r[1] = y; r[2] = z;
}
// r = R v, R row-major 3x3 with row stride `ld`
template <class T> TFN_HD void mulRv(T* r, const T* R, const T* v, int ld = 3) {
  T a = dot3(R[0], R[1], R[2], v[0], v[1], v[2]);
  T b = dot3(R[ld], R[ld + 1], R[ld + 2], v[0], v[1], v[2]);
  T c = dot3(R[2 * ld], R[2 * ld + 1], R[2 * ld + 2], v[0], v[1], v[2]);
  r[0] = a; r[1] = b; r[2] = c;
}
// r = R^T v
template <class T> TFN_HD void mulRtv(T* r, const T* R, const T* v, int ld = 3) {
  T a = dot3(R[0], R[ld], R[2 * ld], v[0], v[1], v[2]);
  T b = dot3(R[1], R[ld + 1], R[2 * ld + 1], v[0], v[1], v[2]);
  T c = dot3(R[2], R[ld + 2], R[2 * ld + 2], v[0], v[1], v[2]);
  r[0] = a; r[1] = b; r[2] = c;
}

// ---------------------------------------------------------------- binary32 transcendental kernels (Cephes)
TFN_HD float f_abs(float x) { return x < 0.0f ? -x : (x == 0.0f ? 0.0f : x); }

TFN_HD float f_pow2i(int n) {  // 2^n, -126 <= n <= 127
  uint32_t u = (uint32_t)(n + 127) << 23;
  float f;
#if defined(__CUDA_ARCH__)
  f = __uint_as_float(u);
#else
  memcpy(&f, &u, 4);
#endif
  return f;
}

TFN_HD float f_atan(float xx) {
  float x = xx < 0.0f ? -xx : xx;
  float y0;
  if (x > 2.414213562373095f) { y0 = 1.5707963267948966f; x = -(1.0f / x); }
  else if (x > 0.4142135623730950f) { y0 = 0.7853981633974483f; x = (x - 1.0f) / (x + 1.0f); }
  else y0 = 0.0f;
  const float z = x * x;
  float p = fma(8.05374449538e-2f, z, -1.38776856032e-1f);
  p = fma(p, z, 1.99777106478e-1f);
  p = fma(p, z, -3.33329491539e-1f);
  float y = fma(p * z, x, x);
  y = y0 + y;
  return xx < 0.0f ? -y : y;
}
TFN_HD float f_atan2(float y, float x) {
  const float PI = 3.14159265358979323846f, PIO2 = 1.5707963267948966f;
  if (x == 0.0f) {
    if (y == 0.0f) return 0.0f;
    return y > 0.0f ? PIO2 : -PIO2;
  }
  if (y == 0.0f) return x > 0.0f ? 0.0f : PI;
  const float z = f_atan(y / x);
  if (x > 0.0f) return z;
  return y > 0.0f ? z + PI : z - PI;
}
TFN_HD float f_asin(float xx) {
  float a = xx < 0.0f ? -xx : xx;
  if (a > 1.0f) a = 1.0f;  // callers pass components of unit vectors; clamp instead of NaN
  if (a < 1.0e-4f) return xx;
  float z, x; bool flag;
  if (a > 0.5f) { z = 0.5f * (1.0f - a); x = sqrtf(z); flag = true; }
  else { x = a; z = x * x; flag = false; }
  float p = fma(4.2163199048e-2f, z, 2.4181311049e-2f);
  p = fma(p, z, 4.5470025998e-2f);
  p = fma(p, z, 7.4953002686e-2f);
  p = fma(p, z, 1.6666752422e-1f);
  float r = fma(p * z, x, x);
  if (flag) { r = r + r; r = 1.5707963267948966f - r; }
  return xx < 0.0f ? -r : r;
}
TFN_HD float f_tan(float xx) {
  float x = xx < 0.0f ? -xx : xx;
  if (x > 8192.0f) return 0.0f;
  int j = (int)(1.27323954473516f * x);
  float y = (float)j;
  if (j & 1) { j += 1; y += 1.0f; }
  float z = ((x - y * 0.78515625f) - y * 2.4187564849853515625e-4f) - y * 3.77489497744594108e-8f;
  const float zz = z * z;
  if (x > 1.0e-4f) {
    float p = fma(9.38540185543e-3f, zz, 3.11992232697e-3f);
    p = fma(p, zz, 2.44301354525e-2f);
    p = fma(p, zz, 5.34112807005e-2f);
    p = fma(p, zz, 1.33387994085e-1f);
    p = fma(p, zz, 3.33331568548e-1f);
    y = fma(p * zz, z, z);
  } else y = z;
  if (j & 2) y = -1.0f / y;
  return xx < 0.0f ? -y : y;
}
TFN_HD float f_exp(float xx) {
  float x = xx;
  if (x > 88.0f) x = 88.0f;
  if (x < -87.0f) x = -87.0f;
  float z = floorf(fma(1.44269504088896341f, x, 0.5f));
  const int n = (int)z;
  x = x - z * 0.693359375f;
  x = x - z * -2.12194440e-4f;
  z = x * x;
  float p = fma(1.9875691500e-4f, x, 1.3981999507e-3f);
  p = fma(p, x, 8.3334519073e-3f);
  p = fma(p, x, 4.1665795894e-2f);
  p = fma(p, x, 1.6666665459e-1f);
  p = fma(p, x, 5.0000001201e-1f);
  z = fma(p, z, x) + 1.0f;
  return z * f_pow2i(n);
}
TFN_HD float f_tanh(float xx) {
  float z = xx < 0.0f ? -xx : xx;
  if (z > 9.0f) return xx > 0.0f ? 1.0f : -1.0f;
  if (z >= 0.625f) {
    const float s = f_exp(z + z);
    z = 1.0f - 2.0f / (s + 1.0f);
    return xx < 0.0f ? -z : z;
  }
  z = xx * xx;
  float p = fma(-5.70498872745e-3f, z, 2.06390887954e-2f);
  p = fma(p, z, -5.37397155531e-2f);
  p = fma(p, z, 1.33314422036e-1f);
  p = fma(p, z, -3.33332819422e-1f);
  return fma(p * z, xx, xx);
}

// ---------------------------------------------------------------- binary64 sin/cos for small angles (|x| <= 2.1 rad)
// Used where an angle is turned into a rotation on the device (the per-episode yaw of the step box,
// src/rover_training_1.cc:84-90): Horner evaluation of the Taylor series, identical on CPU and GPU.
TFN_HD void d_sincos(double x, double* s, double* c) {
  const double z = x * x;
  double ps = -1.0 / 121645100408832000.0;          // -1/19!
  ps = fma(ps, z, 1.0 / 355687428096000.0);         //  1/17!
  ps = fma(ps, z, -1.0 / 1307674368000.0);          // -1/15!
  ps = fma(ps, z, 1.0 / 6227020800.0);              //  1/13!
  ps = fma(ps, z, -1.0 / 39916800.0);               // -1/11!
  ps = fma(ps, z, 1.0 / 362880.0);                  //  1/9!
  ps = fma(ps, z, -1.0 / 5040.0);                   // -1/7!
  ps = fma(ps, z, 1.0 / 120.0);                     //  1/5!
  ps = fma(ps, z, -1.0 / 6.0);                      // -1/3!
  *s = fma(ps * z, x, x);
  double pc = 1.0 / 2432902008176640000.0;          //  1/20!
  pc = fma(pc, z, -1.0 / 6402373705728000.0);       // -1/18!
  pc = fma(pc, z, 1.0 / 20922789888000.0);          //  1/16!
  pc = fma(pc, z, -1.0 / 87178291200.0);            // -1/14!
  pc = fma(pc, z, 1.0 / 479001600.0);               //  1/12!
  pc = fma(pc, z, -1.0 / 3628800.0);                // -1/10!
  pc = fma(pc, z, 1.0 / 40320.0);                   //  1/8!
  pc = fma(pc, z, -1.0 / 720.0);                    // -1/6!
  pc = fma(pc, z, 1.0 / 24.0);                      //  1/4!
  pc = fma(pc, z, -0.5);
  *c = fma(pc, z, 1.0);
}

// ---------------------------------------------------------------- typed front ends
TFN_HD float atan2(float y, float x) { return f_atan2(y, x); }
TFN_HD float asin(float x) { return f_asin(x); }
TFN_HD float tan(float x) { return f_tan(x); }
TFN_HD float tanh(float x) { return f_tanh(x); }
#if !defined(__CUDA_ARCH__)
inline double atan2(double y, double x) { return ::atan2(y, x); }
inline double asin(double x) { return ::asin(x > 1.0 ? 1.0 : (x < -1.0 ? -1.0 : x)); }
inline double tan(double x) { return ::tan(x); }
inline double tanh(double x) { return ::tanh(x); }
#endif

}  // namespace tfn

#endif /* TFODE_NUMERICS_H_ */
