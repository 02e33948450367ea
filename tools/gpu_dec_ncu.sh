#!/bin/bash
# times + key counters of the decode back-end kernels on the small workload (48 streams x 5 s)
mkdir -p gpurun_out
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,lts__t_sectors_op_read.sum,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active
PAC_LIB_NAME=${PAC_LIB_NAME:-libpac_b200.so} python tools/small_codec.py ${1:-192} ${2:-5} > gpurun_out/small_plain.log 2>&1 &&
ncu --metrics $M --clock-control none -k regex:'k_decode|k_ola' -s 3 -c 3 --csv --log-file gpurun_out/ncu_dec.csv env PAC_LIB_NAME=${PAC_LIB_NAME:-libpac_b200.so} python tools/small_codec.py ${1:-192} ${2:-5} > gpurun_out/ncu_dec.log 2>&1
python - <<'PY'
import csv, collections
rows = list(csv.reader(l for l in open("gpurun_out/ncu_dec.csv") if l.startswith('"')))
h = rows[0]; ki = h.index("Kernel Name"); mi = h.index("Metric Name"); vi = h.index("Metric Value"); ii = h.index("ID")
d = collections.OrderedDict()
for r in rows[1:]: d.setdefault((int(r[ii]), r[ki][:40]), {})[r[mi]] = r[vi]
for k in d:
    print(k)
    for m, x in d[k].items():
        print("    %-80s %s" % (m.replace("smsp__average_warps_issue_stalled_", "stall_").replace("_per_issue_active.ratio", ""), x))
PY
