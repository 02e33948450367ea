#!/bin/bash
# bench only (no tests), short: prints stage times
python bench.py --no-cpu --no-mdct --no-fp32 --steps 5 --warmup 3 > gpurun_out/bench_q2.json 2> gpurun_out/bench_q2.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_q2.json").read().strip().splitlines()[-1])
r = d["roofline"]
print("value %.0f e2e %.0f enc %.2f dec %.2f" % (d["value"], d["e2e"]["value"], r["k_encode_ms"], r["k_decode_ms"]))
PY
tail -2 gpurun_out/bench_q2.err
