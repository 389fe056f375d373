#!/bin/bash
# ncu --set full capture of the dominant kernel (trace_persistent) at bounce 0,1,2 of the C3 workload
mkdir -p gpurun_out
timeout 300 python bench.py --quick --overlap 0 --spp 4 --steps 1 --warmup 1 > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:trace_ -c 3 \
    -f -o gpurun_out/${1:-trace_r1} python bench.py --quick --overlap 0 --spp 4 --steps 1 --warmup 1 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
