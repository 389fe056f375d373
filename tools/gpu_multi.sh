#!/bin/bash
# multi-GPU bench lines: tools/gpu_multi.sh N   (run under gpurun --gpus N)
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n$N.log 2>&1; echo "exit $?" >> gpurun_out/bench_n$N.log
tail -2 gpurun_out/bench_n$N.log | cut -c1-400
timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 --scaling strong > gpurun_out/bench_strong_n$N.log 2>&1; tail -1 gpurun_out/bench_strong_n$N.log | cut -c1-300
if [ "$N" = "8" ]; then
  # configs[3]: 10 M triangles, 3840x2160, 256 spp sharded over the 8 GPUs (32 each), one reduce
  timeout 900 $TR bench.py --gpus 8 --workload C4 --spp 32 --steps 3 --warmup 3 > gpurun_out/bench_c4_n8.log 2>&1; echo "exit $?" >> gpurun_out/bench_c4_n8.log
  tail -2 gpurun_out/bench_c4_n8.log | cut -c1-400
  # configs[4]: incoherent stress, 128 spp per GPU
  timeout 900 $TR bench.py --gpus 8 --workload C5 --steps 3 --warmup 3 > gpurun_out/bench_c5_n8.log 2>&1; echo "exit $?" >> gpurun_out/bench_c5_n8.log
  tail -2 gpurun_out/bench_c5_n8.log | cut -c1-400
fi
