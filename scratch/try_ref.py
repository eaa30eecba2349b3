import sys, os, time, tempfile
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo")
import numpy as np
import parity
from qmap_b200 import architecture as ar, circuit as qcirc, routing
from oracle import refapi
ncx = int(sys.argv[1]); kw = eval(sys.argv[2])
R = parity.ref()
n, cm = ar.heavy_hex_127()
csv = ar.synthetic_calibration_csv(n, cm, 127)
d = tempfile.mkdtemp()
open(d + "/a.arch", "w").write(f"{n}\n" + "\n".join(f"{a} {b}" for a, b in sorted(cm)) + "\n")
open(d + "/c.csv", "w").write(csv)
ra = R.arch_load(d + "/a.arch", d + "/c.csv")
circ = qcirc.random_circuit(127, ncx, 3)
c, t = circ.gate_arrays()
t0 = time.time()
rr = R.map_gates(ra, circ.n_qubits, c, t, refapi.make_config(debug=True, **kw))
print("layers", rr.n_layers, "expanded", rr.expanded_nodes, "gen", rr.generated_nodes, "swaps", rr.swaps, "t", time.time() - t0)
