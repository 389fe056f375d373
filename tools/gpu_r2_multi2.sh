#!/bin/bash
# 8-GPU box: C3 at N=4, configs[3] (C4, 10 M tris, 4K, 256 spp over 8 GPUs) and configs[4] (C5, depth 16, 128 spp at 1 and 8 GPUs)
mkdir -p gpurun_out
run() { # N tag args...
  N=$1; tag=$2; shift 2
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
  if [ "$N" = "1" ]; then TR="python"; fi
  timeout 420 $TR bench.py --gpus $N --no-cpu "$@" > gpurun_out/bench_$tag.log 2>&1; echo "exit $?" >> gpurun_out/bench_$tag.log
  grep '^{' gpurun_out/bench_$tag.log | tail -1 > gpurun_out/r2_bench_$tag.json
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/r2_bench_$tag.json"))
    print("$tag", d["n_gpus"], round(d["value"], 1), "Mrays/s", round(d["ms_per_step"], 3), "ms/frame e2e", round(d["e2e"]["value"], 1), "spp_total", d["config"]["spp_total"], "tail", d["config"]["fused_tail_from"])
except Exception as e:
    print("$tag failed", e)
PY
}
run 4 c3_n4 --steps 20 --warmup 3
run 8 c4_n8 --workload C4 --steps 5 --warmup 3
run 1 c5_n1 --workload C5 --steps 3 --warmup 3
run 8 c5_n8 --workload C5 --steps 10 --warmup 3
