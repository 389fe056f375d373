#!/bin/bash
# final single-GPU pass of a round: whole -m gpu suite, smoke, bench line + reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --durations=6 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "passed|failed|FAILED|Error|exit|^[0-9.]+s " gpurun_out/pytest_gpu.log | tail -12
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_full.log 2>&1; echo "bench exit $?" >> gpurun_out/bench_full.log
grep '^{' gpurun_out/bench_full.log | tail -1 > gpurun_out/r2_bench_c3_final.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>&1
grep '^{' gpurun_out/bench_ref.log | tail -1 > gpurun_out/r2_bench_c3_reference.json
python - <<PY
import json
d = json.load(open("gpurun_out/r2_bench_c3_final.json")); r = d["roofline"]; ref = json.load(open("gpurun_out/r2_bench_c3_reference.json"))
print("value", round(d["value"], 1), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), "frac", round(r["frac"], 4), "trace", round(r["trace_mrays_s"], 1),
      "shade frac", round(r["shade"]["frac"], 3), "traffic", r["traffic"], r["traffic_source"], "cpu", d["cpu_baseline"]["value"] if d["cpu_baseline"] else None,
      "ref", round(ref["value"], 2), ref["metric"] == d["metric"], "launches", d["gpu_launches"], d["clocks"])
PY
