#!/usr/bin/env python
"""Per-kernel summary of one bench step from an ncu launch list (csv) + profiles/traffic.json.
usage: launch_summary.py launches.csv out_summary.txt [tag]"""
import csv, collections, json, sys
src, dst = sys.argv[1], sys.argv[2]
tag = sys.argv[3] if len(sys.argv) > 3 else "r01"
rows = list(csv.reader(l for l in open(src) if l.startswith('"')))
h = rows[0]; ki = h.index("Kernel Name"); mi = h.index("Metric Name"); vi = h.index("Metric Value"); ii = h.index("ID")
per = collections.OrderedDict()
for r in rows[1:]:
    per.setdefault(int(r[ii]), {"name": r[ki]})[r[mi]] = float(r[vi].replace(",", ""))
names = [v["name"].split("(")[0].replace("void ", "") for v in per.values()]
idx = [i for i, n in enumerate(names) if n.startswith("k_transient")]
s0, s1 = idx[1], idx[2]
agg = collections.OrderedDict()
for v in list(per.values())[s0:s1]:
    n = v["name"].split("(")[0].replace("void ", "")
    a = agg.setdefault(n, {"n": 0, "t": 0.0, "rd": 0.0, "wr": 0.0, "f64": 0.0, "inst": 0.0, "f64pct": 0.0, "issue": 0.0})
    a["n"] += 1; a["t"] += v.get("gpu__time_duration.sum", 0)
    a["rd"] += v.get("dram__bytes_read.sum", 0); a["wr"] += v.get("dram__bytes_write.sum", 0)
    a["f64"] += v.get("sm__inst_executed_pipe_fp64.sum", 0); a["inst"] += v.get("smsp__inst_executed.sum", 0)
    t = v.get("gpu__time_duration.sum", 0)          # duration-weighted percentages
    a["f64pct"] += t * v.get("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", 0)
    a["issue"] += t * v.get("smsp__issue_active.avg.pct_of_peak_sustained_active", 0)
tot = sum(a["t"] for a in agg.values())
lines = ["one step = launches %d..%d of the capture (%d launches), %.2f ms of kernel time under ncu (serialised, cold cache)"
         % (s0, s1 - 1, s1 - s0, tot / 1e6)]
for n, a in sorted(agg.items(), key=lambda x: -x[1]["t"]):
    lines.append("%-40s n=%3d  total %8.3f ms (%5.1f %%)  dram rd / wr per step %8.1f / %8.1f MB  fp64 pipe %4.1f %%  issue %4.1f %%  fp64 / all warp-inst %6.1f / %7.1f M"
                 % (n[:40], a["n"], a["t"] / 1e6, 100 * a["t"] / tot, a["rd"] / 1e6, a["wr"] / 1e6,
                    a["f64pct"] / max(a["t"], 1), a["issue"] / max(a["t"], 1), a["f64"] / 1e6, a["inst"] / 1e6))
open(dst, "w").write("# ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__inst_executed_pipe_fp64.sum,"
                     "smsp__inst_executed.sum,sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active,"
                     "smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none "
                     "-k regex:'^k_|DeviceScan' -c 400 python bench.py --steps 2 --warmup 1 (side legs off)\n# full list: profiles/%s_launches.csv\n" % tag
                     + "\n".join(lines) + "\n")
print("\n".join(lines))
tr = {"streams": 192, "seconds": 20.0,
      "source": "profiles/%s_launches.csv (ncu, default bench config, sums over the launches of ONE step)" % tag}
for n, a in agg.items():
    b = n.split("<")[0]
    if not b.startswith("k_"):
        continue
    b = "k_encode" if b == "k_encode_front" else b
    tr[b + "_dram_bytes"] = int(a["rd"] + a["wr"])
    if a["f64"]:
        tr[b + "_fp64_warp_inst"] = int(a["f64"])
        tr[b + "_fp64_pipe_pct"] = round(a["f64pct"] / max(a["t"], 1), 2)
        tr[b + "_issue_active_pct"] = round(a["issue"] / max(a["t"], 1), 2)
json.dump(tr, open("profiles/traffic.json", "w"), indent=1)
