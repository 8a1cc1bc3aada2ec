#!/bin/bash
# Developer tool: compile a variant of the library into build/lib<name>.so for same-box A/B timing (tools/ab_bench.py).
#   tools/build_variant.sh <name> [git-rev]     (no rev: the working tree)
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
NAME=$1; REV=$2
T=$(mktemp -d)
mkdir -p $T/tf_ode_b200 $ROOT/build
if [ -n "$REV" ]; then
  (cd $ROOT && git archive $REV include tf_ode_b200/csrc | tar -x -C $T)
else
  cp -r $ROOT/include $T/; cp -r $ROOT/tf_ode_b200/csrc $T/tf_ode_b200/
fi
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC -Xcompiler -ffp-contract=off \
  -shared -cudart static ${TFODE_NVCC_EXTRA} -o $ROOT/build/lib$NAME.so $T/tf_ode_b200/csrc/tfode_capi.cu -lz 2>/dev/null
rm -rf $T
echo "built build/lib$NAME.so"
