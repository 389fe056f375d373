#!/bin/bash
# experiment runner: C3 full frame per trace variant (VARIANTS), serial and two-lane
mkdir -p gpurun_out
out=gpurun_out/exp.log; : > $out
for v in ${VARIANTS:-0}; do
  for ov in 0 1; do
    echo "== variant $v overlap $ov" >> $out
    timeout 200 python bench.py --quick --workload ${WL:-C3} --steps 3 --warmup 3 --variant $v --overlap $ov 2>&1 | tail -1 | cut -c1-120 >> $out
  done
done
cat $out
