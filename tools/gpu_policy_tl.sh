#!/bin/bash
# Phase timelines of policy_tc_kernel on the GPU box.  Build the tool(s) first, here, from the library's own header (see tools/policy_tc_timeline.cu):
#   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -fmad=false -DTFODE_TC_TIMELINE [-DTFODE_TC_LOOPMARKS] [-DTFODE_TC_NSLOT=3] \
#        -I tf_ode_b200/csrc -I include -o tools/policy_tc_timeline_<name> tools/policy_tc_timeline.cu
# Every tools/policy_tc_timeline_<name> found is run, each under its own timeout.
mkdir -p gpurun_out
for f in tools/policy_tc_timeline_*; do v=${f#tools/policy_tc_timeline_}; [ -x $f ] || continue; for m in ${TL_ROWS:-16384}; do
  timeout 60 $f $m > gpurun_out/tl_${v}_$m.txt 2>&1; echo "== $v rows $m"; head -1 gpurun_out/tl_${v}_$m.txt; grep -A15 "mean over" gpurun_out/tl_${v}_$m.txt | tr -s ' ' | paste -sd' ' | cut -c1-600
  grep -A12 "MMA issue loop" gpurun_out/tl_${v}_$m.txt | tail -4
done; done
