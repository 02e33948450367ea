#!/bin/bash
# usage: tools/ncu_kernel.sh <kernel regex> <count> : key counters of the first <count> matching launches of a short bench run
K=${1:-k_dequant}; C=${2:-2}
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,smsp__thread_inst_executed_per_inst_executed.ratio,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio,smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio,sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum,l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum,l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum
ncu --metrics $M --clock-control none -k regex:"$K" -c $C --csv --log-file gpurun_out/ncu_k.csv python bench.py --no-cpu --no-mdct --no-fp32 --steps 1 --warmup 1 > gpurun_out/ncu_k.log 2>&1
python - <<PY
import csv
rows=list(csv.reader(l for l in open("gpurun_out/ncu_k.csv") if l.startswith('"')))
h=rows[0]; ki=h.index("Kernel Name"); mi=h.index("Metric Name"); vi=h.index("Metric Value"); ii=h.index("ID")
d={}
for r in rows[1:]: d.setdefault((int(r[ii]),r[ki][:30]),{})[r[mi]]=r[vi]
for k in sorted(d):
    print(k)
    for m,v in d[k].items(): print("   %-80s %s"%(m,v))
PY
