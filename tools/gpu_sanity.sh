#!/bin/bash
# sanity call after a container restore: parity tests, smoke, quick per-workload lines, full bench + reference arm
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
for w in C1 C2 C5; do
  timeout 300 python bench.py --quick --workload $w --spp 8 --steps 2 --warmup 3 > gpurun_out/quick_$w.log 2>&1
done
timeout 900 python bench.py > gpurun_out/bench_full.log 2>&1; echo "bench exit $?" >> gpurun_out/bench_full.log
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log; tail -n1 gpurun_out/quick_C*.log | cut -c1-400; tail -2 gpurun_out/bench_full.log | cut -c1-900
