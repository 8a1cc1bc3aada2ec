#!/usr/bin/env python
"""Compare two traces (trace_format.h): oracle/libode_harness/trace_oracle vs trace_libode.  TEST INFRASTRUCTURE.

  python compare_traces.py a.trace b.trace
Prints, per checkpoint, the largest position / quaternion / velocity difference and whether the contact counts agree, and the
first step at which the contact counts differ.  With both programs on dWorldQuickStep the expectation is agreement
to rounding until the first contact-count difference -- the oracle's row order inside an island follows its own canonical contact
order (Deviations: D-ORDER), ODE's follows the hash-space traversal, so trajectories are expected to separate only
through the random-reorder stream, slowly."""
import struct
import sys

import numpy as np


def load(path):
    b = open(path, "rb").read()
    magic, ver, nb, ns = struct.unpack("<4siii", b[:16])
    assert magic == b"TFTR" and ver == 1, path
    rec = np.dtype([("step", "<i4"), ("nc", "<i4"), ("body", "<f8", (nb, 13))])
    return np.frombuffer(b, dtype=rec, count=ns, offset=16)


a, b = load(sys.argv[1]), load(sys.argv[2])
n = min(len(a), len(b))
nb = min(a["body"].shape[1], b["body"].shape[1])
diff_nc = np.nonzero(a["nc"][:n] != b["nc"][:n])[0]
print("steps compared: %d, bodies: %d; first contact-count difference at step %s" % (n, nb, a["step"][diff_nc[0]] if len(diff_nc) else "never"))
for k in sorted(set([0, 9, 99, 499, 999, 1999, 4999, n - 1])):
    if k >= n:
        continue
    d = np.abs(a["body"][k][:nb] - b["body"][k][:nb])
    print("step %6d  contacts %2d / %2d   max |dpos| %.3e  |dquat| %.3e  |dlinvel| %.3e  |dangvel| %.3e"
          % (a["step"][k], a["nc"][k], b["nc"][k], d[:, 0:3].max(), d[:, 3:7].max(), d[:, 7:10].max(), d[:, 10:13].max()))
