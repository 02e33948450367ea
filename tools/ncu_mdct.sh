#!/bin/bash
# key counters of the forward / inverse streaming MDCT kernels (fp32), both pipelines
mkdir -p gpurun_out
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio,smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio,smsp__average_warps_issue_stalled_membar_per_issue_active.ratio,smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio,smsp__average_warps_issue_stalled_misc_per_issue_active.ratio,smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio,smsp__average_warps_issue_stalled_selected_per_issue_active.ratio
for v in ${VARIANTS:-0 8}; do
MDCT_F=262144 MDCT_ITERS=1 PAC_BENCH_VARIANT=$v python tools/mdct_bw.py > gpurun_out/mdct_plain_$v.log 2>&1 &&
MDCT_F=262144 MDCT_ITERS=1 PAC_BENCH_VARIANT=$v ncu --metrics $M --clock-control none -k regex:'k_mdct_fwd|k_imdct' -s ${SKIP:-2} -c 2 --csv --log-file gpurun_out/ncu_mdct_$v.csv python tools/mdct_bw.py > gpurun_out/ncu_mdct_$v.log 2>&1
done
python - <<'PY'
import csv, collections
for v in (0, 8):
    try: rows = list(csv.reader(l for l in open(f"gpurun_out/ncu_mdct_{v}.csv") if l.startswith('"')))
    except Exception as e: print(v, e); continue
    h = rows[0]; ki = h.index("Kernel Name"); mi = h.index("Metric Name"); vi = h.index("Metric Value"); ii = h.index("ID")
    d = collections.OrderedDict()
    for r in rows[1:]: d.setdefault((int(r[ii]), r[ki][:42]), {})[r[mi]] = r[vi]
    for k in d:
        print("variant", v, k)
        for m, x in d[k].items():
            m = m.replace("smsp__average_warps_issue_stalled_", "stall_").replace("_per_issue_active.ratio", "").replace(".avg.pct_of_peak_sustained_active", "%")
            print("    %-48s %s" % (m, x))
PY
