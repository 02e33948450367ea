#!/usr/bin/env python
"""Static SASS size per source file / line for one kernel of an object file.
usage: sass_size.py obj.o kernel_substr [top]"""
import subprocess, sys, re, collections, tempfile, os, glob
obj, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, check=True, capture_output=True)
cub = glob.glob(d + "/*.cubin")[0]
out = subprocess.run(["nvdisasm", "-g", "-c", cub], capture_output=True, text=True).stdout
cur = None; infn = False
byfile = collections.Counter(); byline = collections.Counter()
for ln in out.split("\n"):
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
    if m:
        infn = kern in m.group(1); continue
    if not infn: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
        byfile[cur[0] if cur else "?"] += 16; byline[cur] += 16
print("total", sum(byfile.values()))
for f, c in byfile.most_common(): print(f"{c:8d} {f}")
for k, c in byline.most_common(top): print(f"{c:8d} {k}")
