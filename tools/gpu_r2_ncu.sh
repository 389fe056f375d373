#!/bin/bash
# ncu --set full of the trace kernels of one C3 frame (serial lanes), summarised ON the box:
#   tools/gpu_r2_ncu.sh NAME VARIANT TAIL SPP [KERNEL_REGEX] [COUNT]
# writes gpurun_out/NAME_raw.txt (key metrics per launch) and gpurun_out/NAME_blocks_K.txt (basic-block shares)
NAME=${1:-r2_trace}; VAR=${2:-0}; TAIL=${3:--1}; SPP=${4:-16}; KRE=${5:-trace_wq}; CNT=${6:-4}
mkdir -p gpurun_out
CMD="python bench.py --quick --overlap 0 --spp $SPP --spp-per-pass $SPP --steps 1 --warmup 1 --variant $VAR --tail $TAIL"
timeout 300 $CMD > gpurun_out/${NAME}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${NAME}_plain.log; exit 1; }
tail -1 gpurun_out/${NAME}_plain.log | cut -c1-200
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:$KRE -c $CNT -f -o /tmp/$NAME $CMD > gpurun_out/${NAME}_ncu.log 2>&1
tail -2 gpurun_out/${NAME}_ncu.log
ncu -i /tmp/$NAME.ncu-rep --page raw --csv > /tmp/${NAME}_raw.csv 2>/dev/null
python tools/ncu_keys.py /tmp/${NAME}_raw.csv > gpurun_out/${NAME}_raw.txt 2>&1
python - <<PY >> gpurun_out/${NAME}_raw.txt 2>&1
import csv
rows = list(csv.reader(open("/tmp/${NAME}_raw.csv")))
hdr, units, data = rows[0], rows[1], rows[2:]
extra = ["l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
         "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum",
         "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum",
         "l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum", "l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed",
         "smsp__inst_executed_op_shared_atom.sum", "smsp__inst_executed_op_shared_ld.sum", "smsp__inst_executed_op_shared_st.sum",
         "smsp__inst_executed_op_global_ld.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed", "Kernel Name", "launch__grid_size"]
for k in extra:
    if k in hdr:
        i = hdr.index(k)
        print(f"{k:90s} {units[i]:8s}", [r[i][:40] for r in data])
PY
for k in 0 1 2; do python tools/ncu_blocks.py /tmp/$NAME.ncu-rep $k > gpurun_out/${NAME}_blocks_$k.txt 2>&1; done
cut -c1-260 gpurun_out/${NAME}_raw.txt
