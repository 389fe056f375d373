#!/bin/bash
# round 3: ncu --set full of the trace launches of one frame + per-instruction shared-memory conflict attribution
#   tools/gpu_r3_ncu.sh NAME WORKLOAD SPP TAIL [KERNEL_REGEX] [COUNT] [SKIP]
NAME=${1:-r3_trace}; WL=${2:-C3}; SPP=${3:-16}; TAIL=${4:--1}; KRE=${5:-trace_wq}; CNT=${6:-3}; SKIP=${7:-0}
mkdir -p gpurun_out
CMD="python bench.py --quick --workload $WL --overlap 0 --spp $SPP --spp-per-pass $SPP --steps 1 --warmup 1 --tail $TAIL"
timeout 300 $CMD > gpurun_out/${NAME}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${NAME}_plain.log; exit 1; }
tail -1 gpurun_out/${NAME}_plain.log | cut -c1-200
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:$KRE -s $SKIP -c $CNT -f -o /tmp/$NAME $CMD > gpurun_out/${NAME}_ncu.log 2>&1
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
         "l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum", "smsp__inst_executed_op_shared_ld.sum", "smsp__inst_executed_op_shared_st.sum",
         "smsp__inst_executed_op_global_ld.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "Kernel Name", "launch__grid_size"]
for k in extra:
    if k in hdr:
        i = hdr.index(k)
        print(f"{k:90s} {units[i]:8s}", [r[i][:40] for r in data])
PY
for k in $(seq 0 $((CNT-1))); do python tools/ncu_lds_conflicts.py /tmp/$NAME.ncu-rep $k > gpurun_out/${NAME}_lds_$k.txt 2>&1; python tools/ncu_blocks.py /tmp/$NAME.ncu-rep $k > gpurun_out/${NAME}_blocks_$k.txt 2>&1; done
cut -c1-200 gpurun_out/${NAME}_raw.txt | head -60
head -24 gpurun_out/${NAME}_lds_1.txt
