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
    a = agg.setdefault(n, {"n": 0, "t": 0.0, "rd": 0.0, "wr": 0.0})
    a["n"] += 1; a["t"] += v.get("gpu__time_duration.sum", 0)
    a["rd"] += v.get("dram__bytes_read.sum", 0); a["wr"] += v.get("dram__bytes_write.sum", 0)
tot = sum(a["t"] for a in agg.values())
lines = ["one step = launches %d..%d of the capture (%d launches), %.2f ms of kernel time under ncu (serialised, cold cache)"
         % (s0, s1 - 1, s1 - s0, tot / 1e6)]
for n, a in sorted(agg.items(), key=lambda x: -x[1]["t"]):
    lines.append("%-44s n=%3d  total %8.3f ms (%5.1f %%)  avg %7.3f ms  dram rd / wr per launch %8.1f / %8.1f MB"
                 % (n[:44], a["n"], a["t"] / 1e6, 100 * a["t"] / tot, a["t"] / 1e6 / a["n"], a["rd"] / a["n"] / 1e6, a["wr"] / a["n"] / 1e6))
open(dst, "w").write("# ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none "
                     "-k regex:'^k_|DeviceScan' -c 400 python bench.py --steps 2 --warmup 1 --no-cpu\n# full list: profiles/%s_launches.csv\n" % tag
                     + "\n".join(lines) + "\n")
print("\n".join(lines))
tr = {"streams": 192, "seconds": 20.0,
      "source": "profiles/%s_launches.csv (ncu dram__bytes_read.sum + dram__bytes_write.sum, default bench config, average per launch over one step)" % tag}
for n, a in agg.items():
    b = n.split("<")[0]
    if b.startswith("k_"):
        tr[b + "_dram_bytes"] = int((a["rd"] + a["wr"]) / a["n"])
json.dump(tr, open("profiles/traffic.json", "w"), indent=1)
