import sys, os, time, tempfile
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo")
import numpy as np
import parity
from qmap_b200 import architecture as ar, circuit as qcirc
from oracle import refapi
M = int(sys.argv[1]); K = int(sys.argv[2])
R = parity.ref()
n, cm = ar.heavy_hex_127()
a_txt = f"{n}\n" + "\n".join(f"{a} {b}" for a, b in sorted(cm)) + "\n"
d = tempfile.mkdtemp()
open(d + "/a.arch", "w").write(a_txt); open(d + "/c.csv", "w").write(ar.synthetic_calibration_csv(n, cm, 127))
ra = R.arch_load(d + "/a.arch", d + "/c.csv")
kw = dict(heuristic="fidelity_best_location", lookahead_heuristic="none", initial_layout="identity")
mx = []
for s in range(K):
    circ = qcirc.random_circuit(127, M, 1000 + s)
    c, t = circ.gate_arrays()
    rr = R.map_gates(ra, circ.n_qubits, c, t, refapi.make_config(debug=True, **kw))
    g = rr.calls["generated_nodes"]; e = rr.calls["expanded_nodes"]
    mx.append((int(g.max()), int(g.sum()), int(e.sum()), int(rr.calls["solution_depth"].max())))
    print(s, mx[-1], flush=True)
