#!/usr/bin/env python
"""Local-memory (spill / stack) instructions of a kernel by source line.
  python tools/ncu_local.py x.ncu-rep route_hN.o N"""
import re, csv, collections, subprocess, sys, os, tempfile
rep, lib, H = sys.argv[1], sys.argv[2], sys.argv[3]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, stdout=subprocess.DEVNULL)
want = ".text._ZN3qmg12route_kernelILi%sE" % H
for cub in sorted(f for f in os.listdir(tmp) if f.endswith(".cubin")):
    d = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
    if want in d:
        dis = d.split("\n")
        break
start = [i for i, l in enumerate(dis) if want in l and ".section" in l][0]
cur = None; addr2line = {}
pat = re.compile(r"/\*([0-9a-f]{4,})\*/")
for l in dis[start:]:
    if l.startswith("\t.section") and "route_kernel" not in l and addr2line: break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = pat.search(l)
    if m and cur: addr2line[int(m.group(1), 16)] = cur
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.split("\n")))
hdr = rows[1]; data = [r for r in rows[2:] if len(r) == len(hdr)]
ia = hdr.index("Instructions Executed"); iad = hdr.index("Address"); isrc = hdr.index("Source")
base = min(int(r[iad], 16) for r in data)
agg = collections.Counter()
for r in data:
    s = r[isrc]
    if "LDL" in s or "STL" in s:
        key = addr2line.get(int(r[iad], 16) - base)
        agg[(key, "LDL" if "LDL" in s else "STL")] += int(r[ia] or 0)
tot = sum(agg.values())
print("local instructions executed", tot)
src = open(os.environ.get("NCU_SRC", "/root/repo/mqt-qmap_b200/csrc/route_core.hpp")).read().split("\n")
for (k, op), n in agg.most_common(40):
    txt = src[k[1] - 1].strip()[:70] if k and k[0] == "route_core.hpp" else ""
    print(k, op, n, round(100 * n / tot, 1), "|", txt)
