#!/bin/bash
# phase timelines of the policy_tc_kernel variants built by hand (tools/policy_tc_timeline_*), each under its own timeout
mkdir -p gpurun_out
for f in tools/policy_tc_timeline_*; do v=${f#tools/policy_tc_timeline_}; [ -x $f ] || continue; for m in ${TL_ROWS:-16384}; do
  timeout 60 $f $m > gpurun_out/tl_${v}_$m.txt 2>&1; echo "== $v rows $m"; head -1 gpurun_out/tl_${v}_$m.txt; grep -A15 "mean over" gpurun_out/tl_${v}_$m.txt | tr -s ' ' | paste -sd' ' | cut -c1-600
done; done


for f in gpurun_out/tl_*_16384.txt; do echo $f; grep -A12 "MMA issue loop" $f | tail -8; done
