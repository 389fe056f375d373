#!/bin/bash
# quick GPU iteration: parity tests + short C3 bench per trace variant (+ optional extra args)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
for v in ${VARIANTS:-0}; do
  timeout 300 python bench.py --quick --workload C3 --spp 16 --steps 3 --warmup 3 --variant $v $EXTRA 2>&1 | tail -1
done
for w in ${WORKLOADS:-}; do
  timeout 300 python bench.py --quick --workload $w --spp 8 --steps 2 --warmup 3 $EXTRA 2>&1 | tail -1
done
