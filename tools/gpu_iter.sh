#!/bin/bash
# quick GPU iteration: parity tests, then a short bench without the side legs
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -15 gpurun_out/pytest_gpu.log
python bench.py --no-cpu --no-mdct --steps 5 --warmup 3 > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_quick.json").read().strip().splitlines()[-1])
r = d["roofline"]
print("value %.0f e2e %.0f k_enc %.2f k_dec %.2f | fp32 %s | r96 %s" % (d["value"], d["e2e"]["value"], r["k_encode_ms"], r["k_decode_ms"],
      {k: round(v, 2) for k, v in d.get("fp32", {}).items() if k.endswith("_ms")}, {k: round(v, 2) for k, v in d.get("rate96", {}).items() if k.endswith("_ms")}))
PY
tail -3 gpurun_out/bench_quick.err
