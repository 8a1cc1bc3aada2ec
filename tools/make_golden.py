"""Generates tests/golden/*.npy from the REFERENCE'S OWN code, run in this container.

  biquad_x.npy / biquad_y.npy : the reference's filters::LP_second_order_bilinear<double> (Filters/cpp/filters.cc,
      compiled in place by `make -C oracle ref` into oracle/_ref/libref_filters.so), constructed exactly as
      src/rover_1.cc:244 does (Te = 0.001f, w0 = 4 pi, Q = 0.5, in place on one double), driven by a seeded
      input made of a step, a chirp and noise.
Run from the repo root:  make -C oracle ref && python tools/make_golden.py
/root/reference is needed only here; the tests read the committed .npy files.
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    ref = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_filters.so"))
    ref.ref_biquad_run.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    rng = np.random.RandomState(20260411)
    n = 4000
    t = np.arange(n) * 0.001
    x = np.where(t > 0.2, 100.0, 0.0) + 30.0 * np.sin(2 * np.pi * (1 + 20 * t) * t) + 5.0 * rng.randn(n)
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.zeros_like(x)
    ref.ref_biquad_run(x.ctypes.data_as(C.c_void_p), n, y.ctypes.data_as(C.c_void_p))
    os.makedirs(os.path.join(ROOT, "tests", "golden"), exist_ok=True)
    np.save(os.path.join(ROOT, "tests", "golden", "biquad_x.npy"), x)
    np.save(os.path.join(ROOT, "tests", "golden", "biquad_y.npy"), y)
    print("biquad golden: n=%d, y[-1]=%.17g" % (n, y[-1]))


if __name__ == "__main__":
    main()
