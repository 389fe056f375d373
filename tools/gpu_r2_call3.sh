#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "passed|failed|FAILED|Error|exit" gpurun_out/pytest_gpu.log | tail -12
timeout 900 python tools/exp_grid.py C3 "21,-1,1,64;24,-1,1,64;25,-1,1,64;26,-1,1,64;21,-1,0,64;21,0,1,8;21,0,0,8;21,-1,0,8" > gpurun_out/grid2_c3.log 2>&1
cut -c1-300 gpurun_out/grid2_c3.log
