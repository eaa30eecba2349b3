import sys, os, time, tempfile
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo")
import parity
from qmap_b200 import architecture as ar, circuit as qcirc
from oracle import refapi
nq = int(sys.argv[1]); ncx = int(sys.argv[2]); kw = eval(sys.argv[3])
R = parity.ref()
n, cm = ar.heavy_hex_127()
csv = ar.synthetic_calibration_csv(n, cm, 127)
d = tempfile.mkdtemp()
open(d + "/a.arch", "w").write(f"{n}\n" + "\n".join(f"{a} {b}" for a, b in sorted(cm)) + "\n")
open(d + "/c.csv", "w").write(csv)
t0 = time.time()
ra = R.arch_load(d + "/a.arch", d + "/c.csv")
print("arch load", time.time() - t0, flush=True)
circ = qcirc.random_circuit(nq, ncx, 3, p_1q=float(sys.argv[4]) if len(sys.argv) > 4 else 0.5)
c, t = circ.gate_arrays()
t0 = time.time()
rr = R.map_gates(ra, circ.n_qubits, c, t, refapi.make_config(debug=True, **kw))
print("layers", rr.n_layers, "expanded", rr.expanded_nodes, "gen", rr.generated_nodes, "swaps", rr.swaps, "t", time.time() - t0)
