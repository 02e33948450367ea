#!/bin/bash
mkdir -p gpurun_out
for b in "$@"; do
PAC_BENCH_BATCH_CB=$b python bench.py --no-cpu --no-mdct --no-fp32 --no-parity --no-cfg4 --no-copy --steps 5 --warmup 3 > gpurun_out/bench_b$b.json 2> gpurun_out/bench_b$b.err
python - <<PY
import json
d = json.loads(open("gpurun_out/bench_b$b.json").read().strip().splitlines()[-1])
r = d["roofline"]
print("batch $b: value %.0f e2e %.0f enc %.2f dec %.2f" % (d["value"], d["e2e"]["value"], r["k_encode_ms"], r["k_decode_ms"]), {k: round(x, 2) for k, x in r["kernels_ms_per_step"].items()}, "dec-only %.0f" % d["decode_only"]["value"])
PY
tail -n 2 gpurun_out/bench_b$b.err
done
