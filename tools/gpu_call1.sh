#!/bin/bash
# round 2, call 1: all GPU tests, a quick bench, per-phase cycle shares of k_encode
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; tail -25 gpurun_out/pytest_gpu.log
python bench.py --no-cpu --no-mdct --no-fp32 --steps 5 --warmup 3 > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_q.json").read().strip().splitlines()[-1])
r = d["roofline"]
print("value %.0f e2e %.0f enc %.2f dec %.2f" % (d["value"], d["e2e"]["value"], r["k_encode_ms"], r["k_decode_ms"]), r["kernels_ms_per_step"])
PY
tail -2 gpurun_out/bench_q.err
PAC_LIB_NAME=libpac_phase.so python tools/phase_timing.py > gpurun_out/phase.txt 2>&1; cat gpurun_out/phase.txt
