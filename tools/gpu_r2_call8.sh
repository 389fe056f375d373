#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_checked.py tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -4
for b in 6 8; do
  WPT_SHADE_BLOCKS=$b timeout 300 python tools/exp_grid.py C3 "0,-1,1,64;0,-1,0,64;0,0,1,8" 2>&1 | cut -c1-300 | sed "s/^/shade_blocks=$b /"
done
