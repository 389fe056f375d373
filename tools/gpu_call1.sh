#!/bin/bash
# first GPU call of the round: parity tests, trace-variant timings, the full bench line, launch list
mkdir -p gpurun_out
nvidia-smi > gpurun_out/nvidia_smi.txt 2>&1
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
for v in 0 1 2; do
  timeout 300 python bench.py --quick --workload C3 --spp 16 --steps 3 --warmup 3 --variant $v > gpurun_out/quick_v$v.log 2>&1
done
for w in C1 C2 C5; do
  timeout 300 python bench.py --quick --workload $w --spp 8 --steps 2 --warmup 3 > gpurun_out/quick_$w.log 2>&1
done
timeout 600 python bench.py > gpurun_out/bench_full.log 2>&1; echo "bench exit $?" >> gpurun_out/bench_full.log
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1
timeout 300 python bench.py --quick --spp 4 --steps 1 --warmup 1 > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv \
    python bench.py --quick --spp 4 --steps 1 --warmup 1 > gpurun_out/ncu_list.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/smoke.log | tail -2; cat gpurun_out/quick_v*.log; tail -2 gpurun_out/bench_full.log
