#!/bin/bash
# full evidence run: GPU tests, the default bench line, the reference arm, then the ncu launch list of one step
TAG=${1:-r02a}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; tail -n 4 gpurun_out/${TAG}_pytest_gpu.log
python bench.py > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_n1.err; echo "bench rc $?"; tail -n 3 gpurun_out/${TAG}_bench_n1.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference_arm.json 2>/dev/null
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__inst_executed_pipe_fp64.sum,smsp__inst_executed.sum,sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active
CMD="python bench.py --steps 2 --warmup 1 --no-cpu --no-mdct --no-fp32 --no-parity --no-cfg4 --no-copy"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics $M --clock-control none -k regex:'^k_|DeviceScan' -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/ncu.log 2>&1
echo "ncu rc $?"
python - <<PY
import json
d = json.loads(open("gpurun_out/${TAG}_bench_n1.json").read().strip().splitlines()[-1])
r = d["roofline"]
print("value %.0f e2e %.0f" % (d["value"], d["e2e"]["value"]), {k: round(x, 2) for k, x in r["kernels_ms_per_step"].items()})
for k in ("decode_only", "copy_roofline_gbs", "parity", "cfg4"):
    print(k, json.dumps(d.get(k))[:600])
print("mdct", json.dumps(d.get("mdct"))[:900])
PY
