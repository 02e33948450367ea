#!/usr/bin/env python
"""Summarise an ncu report per CUDA source line: samples and instructions.
usage: ncu_lines.py report.ncu-rep kernel_regex [top_n]"""
import csv, subprocess, sys, collections, io
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass",
                      "--kernel-name", "regex:" + kern], capture_output=True, text=True).stdout
fname = None
agg = collections.defaultdict(lambda: [0, 0, ""])   # (file,line) -> samples, inst, text
stall = collections.defaultdict(lambda: collections.Counter())
hdr = None
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] in ("File Name", "File Path"):
        fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No":
        hdr = r; continue
    if r[0] == "Kernel Name" or hdr is None or len(r) < len(hdr):
        continue
    try:
        ln = int(r[0])
    except ValueError:
        continue
    si, ii = hdr.index("# Samples"), hdr.index("Instructions Executed")
    key = (fname, ln)
    try:
        s = int(r[si] or 0); n = int(r[ii] or 0)
    except ValueError:
        continue
    # rows with an Address are SASS rows attributed to this line; line row itself carries totals
    if r[2] == "-":     # cuda line summary row
        agg[key][0] = s; agg[key][1] = n; agg[key][2] = r[1].strip()[:90]
        for j, h in enumerate(hdr):
            if h.startswith("stall_") and "Not Issued" not in h and r[j] not in ("", "0"):
                try: stall[key][h] += int(r[j])
                except ValueError: pass
ts = sum(v[0] for v in agg.values()) or 1
ti = sum(v[1] for v in agg.values()) or 1
print(f"total samples {ts}  total warp-instructions {ti}")
byfile = collections.Counter()
for (f, l), v in agg.items():
    byfile[f] += v[0]
print("by file:", {f: round(c / ts, 3) for f, c in byfile.most_common()})
byfile_i = collections.Counter()
for (f, l), v in agg.items():
    byfile_i[f] += v[1]
print("inst by file:", {f: round(c / ti, 3) for f, c in byfile_i.most_common()})
if len(sys.argv) > 4:      # inst-sorted listing
    for (f, l), v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[4])]:
        print(f"I {f}:{l:<4d} inst {v[1]/ti:6.3f} smp {v[0]/ts:6.3f} {v[2]}")
for (f, l), v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    st = ",".join(f"{k[6:]}:{c}" for k, c in stall[(f, l)].most_common(3))
    print(f"{f}:{l:<4d} smp {v[0]/ts:6.3f} inst {v[1]/ti:6.3f} [{st}] {v[2]}")
