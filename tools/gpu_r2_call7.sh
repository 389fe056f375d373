#!/bin/bash
# round 2: ncu of the stream kernels, launch list, DRAM traffic of the trace launches, sanitizers
mkdir -p gpurun_out
python -m pytest tests/test_shim_main.py -m gpu -q 2>&1 | tail -2
# (1) stream kernels, --set full
CMD="python bench.py --quick --overlap 0 --spp 16 --spp-per-pass 16 --steps 1 --warmup 1 --tail -1"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"shade_kernel|raygen_kernel|film_scale" -c 5 -f -o /tmp/r2s $CMD > gpurun_out/r2s_ncu.log 2>&1
ncu -i /tmp/r2s.ncu-rep --page raw --csv > /tmp/r2s_raw.csv 2>/dev/null
python tools/ncu_keys.py /tmp/r2s_raw.csv > gpurun_out/r2_stream_kernels_ncu.txt 2>&1
python - <<PY >> gpurun_out/r2_stream_kernels_ncu.txt 2>&1
import csv
rows = list(csv.reader(open("/tmp/r2s_raw.csv"))); hdr, units, data = rows[0], rows[1], rows[2:]
for k in ["Kernel Name", "dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum.per_second", "dram__bytes_write.sum.per_second",
          "l1tex__average_t_sectors_per_request_pipe_lsu_mem_global_op_ld.ratio", "l1tex__average_t_sectors_per_request_pipe_lsu_mem_global_op_st.ratio",
          "lts__t_sectors_srcunit_tex_op_read_lookup_hit.sum", "lts__t_sectors_srcunit_tex_op_read_lookup_miss.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
          "gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed", "launch__grid_size", "launch__occupancy_limit_registers"]:
    if k in hdr:
        i = hdr.index(k); print(f"{k:90s} {units[i]:10s}", [r[i][:28] for r in data])
PY
cut -c1-230 gpurun_out/r2_stream_kernels_ncu.txt | head -50
# (2) launch list of the default frame (two lanes, 64 spp)
timeout 300 python bench.py --quick --steps 1 --warmup 1 > gpurun_out/plain64.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2_launches_c3_64spp.csv python bench.py --quick --steps 1 --warmup 1 > gpurun_out/ncu_list.log 2>&1
# (3) DRAM bytes of every trace launch of one serial 16-spp frame
timeout 300 python tools/traffic_frame.py 16 > gpurun_out/traffic_plain.json 2>gpurun_out/traffic_plain.err && \
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:trace_wq --csv --log-file gpurun_out/traffic_ncu.csv python tools/traffic_frame.py 16 > gpurun_out/traffic_ncu.log 2>&1
python - <<PY
import csv, json
live = json.loads(open("gpurun_out/traffic_plain.json").read().strip().splitlines()[-1])
rows = [r for r in csv.reader(open("gpurun_out/traffic_ncu.csv")) if len(r) > 10 and r[0].isdigit()]
per = {}
for r in rows:
    kid, name, metric, unit, val = r[0], r[4], r[-3], r[-2], float(r[-1].replace(",", ""))
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
    per.setdefault(kid, {"name": name[:60]})[metric] = val * scale
ks = sorted(per, key=int)
n_launch = len(ks) // 2            # the script renders the frame twice
frame = ks[n_launch:]
byts = [per[k].get("dram__bytes_read.sum", 0) + per[k].get("dram__bytes_write.sum", 0) for k in frame]
tot = sum(byts)
out = {"source": "ncu dram__bytes_read.sum + dram__bytes_write.sum, every trace_wq launch of one serial C3 frame (tools/traffic_frame.py)",
       "spp": live["spp"], "segments": live["segments"], "live_per_bounce": live["live"], "trace_launches": len(frame),
       "dram_bytes_per_launch": byts, "dram_bytes_total": tot, "dram_bytes_per_ray_weighted": tot / live["segments"]}
json.dump(out, open("gpurun_out/r2_trace_traffic.json", "w"), indent=1)
print("traffic", len(frame), "launches", round(tot / live["segments"], 2), "B/ray")
PY
# (4) sanitizers
bash tools/sanitize.sh 2>&1 | tail -16
