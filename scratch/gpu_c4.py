"""GPU experiment: config 4 (heavy-hex 127, fidelity-aware) batch throughput + parity."""
import sys, os, time, tempfile, ctypes as C
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/mqt-qmap_b200")
import numpy as np
import parity
from qmap_b200 import architecture as ar, circuit as qcirc, routing, _abi
B = int(sys.argv[1]); M = int(sys.argv[2]); NCHK = int(sys.argv[3]) if len(sys.argv) > 3 else 2
spsm = int(sys.argv[4]) if len(sys.argv) > 4 else 0
pool = int(sys.argv[5]) if len(sys.argv) > 5 else 0
maxs = int(sys.argv[6]) if len(sys.argv) > 6 else 0
kw = dict(heuristic="fidelity_best_location", lookahead_heuristic="none", initial_layout="identity")
L = _abi.load()
n, cm = ar.heavy_hex_127()
a = ar.Architecture(n, cm)
d = tempfile.mkdtemp()
open(d + "/a.arch", "w").write(a.to_arch_text()); open(d + "/c.csv", "w").write(ar.synthetic_calibration_csv(n, cm, 127))
t0 = time.time(); a.load_properties(d + "/c.csv"); h = a.handle; print("gpu arch+tables s", time.time() - t0, flush=True)
cfg = routing.make_config(**kw)
circs = [qcirc.random_circuit(127, M, 1000 + s) for s in range(B)]
items = [parity.item_for(L, n, c, cfg) for c in circs]
opts = _abi.QmgRouteOptions(); opts.lean_outputs = 1; opts.searches_per_sm = spsm; opts.pool_nodes = pool; opts.max_searches = maxs
be = routing.CudaBackend(L, h)
for rep in range(int(os.environ.get("REPS", "3"))):
    bufs = routing.pack(n, items, lean=True)
    st = _abi.QmgRouteStats()
    t0 = time.time()
    _abi.check(L, L.qmg_route_batch(h, C.byref(cfg), len(items), C.addressof(bufs.inst), C.addressof(bufs.res), C.byref(opts), C.byref(st)))
    dt = time.time() - t0
    outs = routing.unpack(bufs)
    bad = [o.status for o in outs if o.status != 0]
    exp = sum(o.expanded_nodes for o in outs); gen = sum(o.generated_nodes for o in outs)
    print(f"B={B} M={M} bad={len(bad)} {bad[:3]} exp={exp} gen={gen} kernel_ms={st.kernel_ms:.2f} e2e_ms={dt*1e3:.1f} nodes/s={exp/st.kernel_ms*1e3:.3e} grid={st.grid} retried={st.retried_instances} launches={st.kernel_launches} maxexp={max(o.expanded_nodes for o in outs)}", flush=True)
if NCHK:
    R = parity.ref(); t0 = time.time(); ra = R.arch_load(d + "/a.arch", d + "/c.csv"); print("ref arch s", time.time() - t0, flush=True)
    outs = routing.route_batch(be, n, routing.make_config(debug=True, **kw), items[:NCHK], want_layer_qubits=True)
    tot = 0; t0 = time.time()
    for c, o in zip(circs[:NCHK], outs):
        cc, tt = c.gate_arrays()
        rr = R.map_gates(ra, c.n_qubits, cc, tt, parity.refapi.make_config(debug=True, **kw))
        tot += rr.expanded_nodes
        print("parity:", repr(parity.compare(o, rr)), rr.expanded_nodes, flush=True)
    dt = time.time() - t0
    print(f"ref 1 core: {tot/dt:.1f} nodes/s ({tot} nodes, {dt:.1f}s)")
