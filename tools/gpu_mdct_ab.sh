#!/bin/bash
# median-of-7 MDCT leg of the bench for each library given
mkdir -p gpurun_out
for l in "$@"; do
PAC_LIB_NAME=$l python bench.py --no-cpu --no-fp32 --no-parity --no-cfg4 --no-copy --steps 2 --warmup 1 > gpurun_out/mdct_$l.json 2> gpurun_out/mdct_$l.err
python - <<PY
import json
d = json.loads(open("gpurun_out/mdct_$l.json").read().strip().splitlines()[-1])["mdct"]
for k in ("f32", "f64"):
    m = d[k]
    print("$l", k, "fwd %.3f ms %.1f%%  inv %.3f ms %.1f%%  kaiser fwd %.1f%% inv %.1f%%" % (m["fwd_ms"], 100 * m["fwd_frac"], m["inv_ms"], 100 * m["inv_frac"], 100 * m.get("kaiser_fwd_frac", 0), 100 * m.get("kaiser_inv_frac", 0)), m.get("fwd_ms_all"))
PY
done
