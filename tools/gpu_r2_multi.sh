#!/bin/bash
# multi-GPU call: tools/gpu_r2_multi.sh N  (under gpurun --gpus N)
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
[ -n "$SKIP_TESTS" ] || timeout 600 python -m pytest tests/test_gpu_dist.py tests/test_shim_main.py tests/test_gpu_round2.py -m gpu -q -k "two or film_reduce or interaction" > gpurun_out/pytest_multi_n$N.log 2>&1
tail -4 gpurun_out/pytest_multi_n$N.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/bench_n1.log 2>&1; tail -1 gpurun_out/bench_n1.log | cut -c1-160
CONFIGS=("" "--depth-shard 0" "--tail -1")
if [ -n "$ONLY_DEFAULT" ]; then CONFIGS=("" "--depth-shard 0"); fi
for extra in "${CONFIGS[@]}"; do
  tag=$(echo "$extra" | tr -d ' -'); tag=${tag:-default}
  timeout 150 $TR bench.py --gpus $N --steps 20 --warmup 3 $extra > gpurun_out/bench_n${N}_$tag.log 2>&1; echo "exit $?" >> gpurun_out/bench_n${N}_$tag.log
  tail -2 gpurun_out/bench_n${N}_$tag.log | cut -c1-200
done
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/bench_n*_*.log")) + ["gpurun_out/bench_n1.log"]:
    for ln in open(f):
        if ln.startswith("{"):
            d = json.loads(ln)
            print(f, d["n_gpus"], round(d["value"], 1), "ms/frame", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), round(d["e2e"]["ms_per_frame"], 3), d["config"]["spp_total"], d["config"].get("fused_tail_from"))
PY
