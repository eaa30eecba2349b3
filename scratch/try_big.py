import sys, os, time, tempfile
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo")
import numpy as np
import parity
from qmap_b200 import architecture as ar, circuit as qcirc, routing
from emu_backend import EmuBackend
import emu_backend
from oracle import refapi

which = sys.argv[1]
ncx = int(sys.argv[2]) if len(sys.argv) > 2 else 200
R = parity.ref()
if which == "hh27":
    n, cm = ar.heavy_hex_27()
    ra = R.arch_create(n, sorted(cm))
    kw = dict()
    circ = qcirc.random_circuit(20, ncx, 3)
elif which == "hh127":
    n, cm = ar.heavy_hex_127()
    csv = ar.synthetic_calibration_csv(n, cm, 127)
    d = tempfile.mkdtemp()
    open(d + "/a.arch", "w").write(f"{n}\n" + "\n".join(f"{a} {b}" for a, b in sorted(cm)) + "\n")
    open(d + "/c.csv", "w").write(csv)
    ra = R.arch_load(d + "/a.arch", d + "/c.csv")
    kw = dict(heuristic="fidelity_best_location", lookahead_heuristic="none", layering="disjoint_2q_blocks") if len(sys.argv) < 4 else eval(sys.argv[3])
    circ = qcirc.random_circuit(127, ncx, 3)
lib = emu_backend.lib()
be = EmuBackend.from_ref_arch(ra)
t = time.time()
msg, out, rr = parity.run_pair(be, lib, ra, circ, **kw)
print("msg:", repr(msg), "layers", rr.n_layers, "expanded", rr.expanded_nodes, "gen", rr.generated_nodes, "swaps", rr.swaps, "status", out.status, "t", time.time() - t)
