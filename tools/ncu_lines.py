"""Attribute ncu per-SASS-instruction counts to source lines (needs -lineinfo).
usage: ncu_lines.py <report.ncu-rep> <lib.so> [topN]"""
import re, csv, collections, subprocess, sys, os, tempfile
rep, lib = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, stdout=subprocess.DEVNULL)
want = ".text._ZN3qmg12route_kernelILi%sE" % (sys.argv[4] if len(sys.argv) > 4 else "0")
dis = None
# (one cubin per translation unit: take the one that holds the kernel)
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
ia = hdr.index("Instructions Executed"); iad = hdr.index("Address"); isamp = hdr.index("# Samples")
ist = {k: hdr.index(k) for k in ("stall_no_inst", "stall_long_sb", "stall_wait", "stall_short_sb", "stall_branch_resolving")}
base = min(int(r[iad], 16) for r in data)
agg = collections.Counter(); sagg = collections.Counter(); static = collections.Counter()
stl = collections.defaultdict(collections.Counter)
for r in data:
    key = addr2line.get(int(r[iad], 16) - base)
    n = int(r[ia] or 0)
    agg[key] += n; sagg[key] += int(r[isamp] or 0); static[key] += 1
    for k, i in ist.items(): stl[key][k] += int(r[i] or 0)
tot = sum(agg.values()); stot = sum(sagg.values())
print("total inst", tot, "static", len(data), "samples", stot)
# the source the capture was built from (NCU_SRC=path; default: the tree's)
src = open(os.environ.get("NCU_SRC", "/root/repo/mqt-qmap_b200/csrc/route_core.hpp")).read().split("\n")
for k, n in sagg.most_common(top):
    txt = src[k[1] - 1].strip()[:60] if k and k[0] == "route_core.hpp" else ""
    print(k, "samp%", round(100 * n / stot, 1), "inst%", round(100 * agg[k] / tot, 1), "static", static[k], dict(stl[k].most_common(2)), "|", txt)

# per-function aggregation (enclosing QMG_HD definition of each line)
import bisect
defs = [(i + 1, re.sub(r".*?([A-Za-z0-9_]+)\(.*", r"\1", l)) for i, l in enumerate(src) if re.match(r"\s*QMG_(HD|D) ", l)]
dl = [d[0] for d in defs]
fa = collections.Counter(); fs = collections.Counter(); fst = collections.defaultdict(collections.Counter)
for k, nI in agg.items():
    if k and k[0] == "route_core.hpp":
        j = bisect.bisect_right(dl, k[1]) - 1
        name = defs[j][1] if j >= 0 else "?"
    else:
        name = str(k[0]) if k else "none"
    fa[name] += nI; fs[name] += sagg[k]
    for kk, vv in stl[k].items(): fst[name][kk] += vv
print("---- per function")
for name, nI in fa.most_common(30):
    print(f"{name:24s} inst% {100*nI/tot:5.1f}  samp% {100*fs[name]/stot:5.1f}", dict(fst[name].most_common(3)))
# hot code footprint per function: static instructions executed at least once
ff = collections.Counter(); fh = collections.Counter()
thr = tot / 1e5
for r in data:
    key = addr2line.get(int(r[iad], 16) - base)
    nI = int(r[ia] or 0)
    if key and key[0] == "route_core.hpp":
        j = bisect.bisect_right(dl, key[1]) - 1
        name = defs[j][1] if j >= 0 else "?"
    else:
        name = str(key[0]) if key else "none"
    if nI > 0: ff[name] += 1
    if nI > thr: fh[name] += 1
print("---- executed static instr per function (touched / hot>1e-5 of total)")
for name, c in ff.most_common(30):
    print(f"{name:24s} touched {c:5d} ({c*16/1024:5.1f} KB)  hot {fh[name]:5d}")
print("total touched", sum(ff.values()), "hot", sum(fh.values()))
