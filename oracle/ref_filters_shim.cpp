// CPU ORACLE -- TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the reference's own filters::LP_second_order_bilinear<double>
// (Filters/cpp/filters.hh, compiled from /root/reference by `make ref`).  Used to pin the
// biquad restatement in rover_oracle.hpp and to generate tests/golden/biquad_*.npy.
#include "filters.hh"
extern "C" {
// Same construction as src/rover_1.cc:244 : Te=0.001 (float), w0=4*pi, Q=0.5, in-place on one double.
void ref_biquad_run(const double* x, int n, double* y) {
  double cell = 0;
  filters::LP_second_order_bilinear<double> f(0.001, 4 * 3.14159265358979323846, 0.5, &cell, &cell);
  filters::LP_filter<double>& base = f;  // update() (no argument) is hidden in the derived class
  for (int i = 0; i < n; i++) { cell = x[i]; base.update(); y[i] = cell; }
}
}
