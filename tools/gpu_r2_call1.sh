#!/bin/bash
# round 2, first GPU call: the whole -m gpu suite, smoke, the experiment grid (flush scheme x refill threshold x fused tail),
# the bench line and the reference arm
mkdir -p gpurun_out
nvidia-smi > gpurun_out/nvidia_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x --durations=15 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log; tail -2 gpurun_out/smoke.log
timeout 900 python tools/exp_grid.py C3 "20,-1,1,64;21,-1,1,64;22,-1,1,64;23,-1,1,64;21,1,1,64;21,2,1,64;21,3,1,64;21,4,1,64;22,2,1,64;20,2,1,64;21,-1,1,8;21,1,1,8;21,2,1,8;21,3,1,8;20,-1,1,8" > gpurun_out/grid_c3.log 2>&1
cat gpurun_out/grid_c3.log | cut -c1-330
timeout 600 python bench.py > gpurun_out/bench_full.log 2>&1; echo "bench exit $?" >> gpurun_out/bench_full.log
tail -2 gpurun_out/bench_full.log | cut -c1-1500
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1; tail -1 gpurun_out/bench_ref.log | cut -c1-400
