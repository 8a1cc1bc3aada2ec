#!/bin/bash
# Developer tool: A/B of library variants (build/lib<name>.so) against the in-tree library on every workload and both episode segments.
cp tf_ode_b200/csrc/libtfode_b200.so build/libbase.so
echo "== config2 cheap"; python tools/ab_bench.py --rounds 1 --workload config2 base "$@"
echo "== config2 groove"; python tools/ab_bench.py --rounds 1 --workload config2 --pre 42 --steps 6 base "$@"
echo "== config3 early"; python tools/ab_bench.py --rounds 1 --workload config3 --pre 10 --steps 6 base "$@"
echo "== config3 late";  python tools/ab_bench.py --rounds 1 --workload config3 --pre 60 --steps 6 base "$@"
echo "== config4"; python tools/ab_bench.py --rounds 1 --workload config4 base "$@"
echo "== config4 late"; python tools/ab_bench.py --rounds 1 --workload config4 --pre 60 --steps 6 base "$@"
echo "== config5"; python tools/ab_bench.py --rounds 1 --workload config5 --steps 3 base "$@"
echo "== config5 late"; python tools/ab_bench.py --rounds 1 --workload config5 --pre 15 --steps 3 base "$@"
