#!/bin/bash
# A/B: new front end (variant 0) against the previous one (variant 1); phase shares of the new one
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -8 gpurun_out/pytest_gpu.log
for v in 0 1 3; do
PAC_BENCH_VARIANT=$v python bench.py --no-cpu --no-mdct --no-fp32 --steps 5 --warmup 3 > gpurun_out/bench_v$v.json 2> gpurun_out/bench_v$v.err
python - <<PY
import json
d = json.loads(open("gpurun_out/bench_v$v.json").read().strip().splitlines()[-1])
r = d["roofline"]
print("variant $v: value %.0f e2e %.0f enc %.2f dec %.2f" % (d["value"], d["e2e"]["value"], r["k_encode_ms"], r["k_decode_ms"]), r["kernels_ms_per_step"])
PY
tail -2 gpurun_out/bench_v$v.err
done
PAC_LIB_NAME=libpac_phase.so python tools/phase_timing.py > gpurun_out/phase.txt 2>&1; cat gpurun_out/phase.txt
