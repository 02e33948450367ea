#!/bin/bash
# A/B of library variants: tools/gpu_libs.sh libA.so libB.so ...  (short bench per library, kernel times per step)
mkdir -p gpurun_out
for l in "$@"; do
PAC_LIB_NAME=$l python bench.py --no-cpu --no-mdct --no-fp32 --no-parity --no-cfg4 --no-copy --steps 6 --warmup 3 > gpurun_out/bench_$l.json 2> gpurun_out/bench_$l.err
python - <<PY
import json
d = json.loads(open("gpurun_out/bench_$l.json").read().strip().splitlines()[-1])
r = d["roofline"]
print("$l: value %.0f e2e %.0f dec-only %.0f" % (d["value"], d["e2e"]["value"], d["decode_only"]["value"]), {k: round(x, 2) for k, x in r["kernels_ms_per_step"].items()})
PY
tail -n 2 gpurun_out/bench_$l.err
done
