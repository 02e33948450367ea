#!/bin/bash
mkdir -p gpurun_out
for b in "$@"; do
PAC_BENCH_HOST_CHUNK_MB=$b python bench.py --no-cpu --no-mdct --no-fp32 --no-parity --no-cfg4 --no-copy --steps 6 --warmup 3 > gpurun_out/bench_c$b.json 2> gpurun_out/bench_c$b.err
python - <<PY
import json
d = json.loads(open("gpurun_out/bench_c$b.json").read().strip().splitlines()[-1])
r = d["roofline"]
print("chunk_mb $b: value %.0f e2e %.0f (%.3f) dec-only %.0f e2e %.0f" % (d["value"], d["e2e"]["value"], d["e2e"]["value"] / d["value"], d["decode_only"]["value"], d["decode_only"]["e2e"]["value"]), {k: round(x, 2) for k, x in r["kernels_ms_per_step"].items()})
PY
tail -n 2 gpurun_out/bench_c$b.err
done
