#!/usr/bin/env python
"""Copy the evidence of the last tools/gpu_round.sh run from gpurun_out/ into profiles/ (tracked):
usage: refresh_profiles.py <ncu-rep name without extension> <kernel mangled-name substring> "<title>" """
import csv, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.chdir(ROOT)
rep, ksub, title = sys.argv[1], sys.argv[2], sys.argv[3]
raw = subprocess.run(["ncu", "-i", f"gpurun_out/{rep}.ncu-rep", "--page", "raw", "--csv"], capture_output=True, text=True).stdout
tmp = "smallpt-enoki-optix_b200/build/scratch/raw.csv"
os.makedirs(os.path.dirname(tmp), exist_ok=True)
open(tmp, "w").write(raw)
keys = subprocess.run([sys.executable, "tools/ncu_keys.py", tmp], capture_output=True, text=True).stdout
blocks = subprocess.run([sys.executable, "tools/ncu_blocks.py", f"gpurun_out/{rep}.ncu-rep", "0"], capture_output=True, text=True).stdout
open("profiles/r1_trace_wq_ncu.txt", "w").write(f"# ncu --set full, {title}, C3 4 spp in one pass (--overlap 0): bounce 0, 1, 2 (gpurun_out/{rep}.ncu-rep)\n{keys}\n## basic-block shares, bounce 0\n{blocks}")
rows = list(csv.reader(raw.splitlines()))
hdr, data = rows[0], rows[2:]
col = lambda k: [float(r[hdr.index(k)].replace(",", "")) for r in data]
rd, wr = col("dram__bytes_read.sum"), col("dram__bytes_write.sum")
unit = rows[1][hdr.index("dram__bytes_read.sum")]
mul = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}[unit]
rays = {"bounce0": 8294400, "bounce1": 7479372, "bounce2": 2617427}
per = {k: round((rd[i] + wr[i]) * mul / n, 1) for i, (k, n) in enumerate(rays.items())}
w = round(sum((rd[i] + wr[i]) * mul for i in range(3)) / sum(rays.values()), 1)
json.dump({"kernel": "trace_wq", "source": f"ncu --set full, gpurun_out/{rep}.ncu-rep (C3, 4 spp in one pass, --overlap 0; summary in profiles/r1_trace_wq_ncu.txt)",
           "dram_bytes_per_ray": per, "dram_bytes_per_ray_weighted": w, "rays": rays,
           "note": "dram__bytes_read.sum + dram__bytes_write.sum per launch divided by the rays of that launch; algorithmic bytes are 688 per ray: the BVH is L2 resident, HBM traffic is the ray/hit streams"},
          open("profiles/r1_trace_traffic.json", "w"), indent=1)
for src, dst in (("bench_full.log", "r1_bench_c3.json"), ("bench_ref.log", "r1_bench_reference_arm.json")):
    l = [x for x in open(f"gpurun_out/{src}") if x.startswith("{")][-1]
    open(f"profiles/{dst}", "w").write(l)
open("profiles/r1_launches_c3_spp4.csv", "w").write(open("gpurun_out/launches.csv").read())
obj = "smallpt-enoki-optix_b200/build/kernels.o"
out = f"# cuobjdump -sass of the default trace kernel ({title}, sm_100a), opcode histogram then the listing\n"
out += f"# made by: python tools/sass_stats.py {obj} {ksub} --dump\n"
out += subprocess.run([sys.executable, "tools/sass_stats.py", obj, ksub, "--dump"], capture_output=True, text=True).stdout
out += "# shade_kernel / raygen_kernel SASS opcode histograms\n"
for k in ("shade_kernel", "raygen_kernel"):
    out += subprocess.run([sys.executable, "tools/sass_stats.py", obj, k], capture_output=True, text=True).stdout
open("profiles/r1_trace_wq_sass.txt", "w").write(out)
d = json.loads(open("profiles/r1_bench_c3.json").read())
r = d["roofline"]
print("value", d["value"], "ms", d["ms_per_step"], "e2e", d["e2e"]["value"], "frac", r["frac"], "trace", r["trace_mrays_s"], r["share_of_step"], "traffic/ray", per, w)
print(keys[:1800])
