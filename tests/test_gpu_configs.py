"""-m gpu: BASELINE.json's configs 3 and 4 through the C ABI against the
compiled reference (oracle/_ref), call by call.

config 3: the circuit suite on a 27-qubit heavy-hex architecture (batched)
config 4: random 127-qubit circuits on a 127-qubit heavy-hex architecture
          with calibration data, fidelity_best_location; searches here reach
          10^4..10^5 open nodes per layer, which exercises the blocked open
          list (route_core.hpp kBlk) and, with a small first pool, the
          re-routing of instances whose node pool ran out."""
import os

import numpy as np
import pytest

import parity
from qmap_b200 import _abi, routing, workloads
from qmap_b200 import architecture as ar

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def c4(tmp_path_factory):
    d = str(tmp_path_factory.mktemp("c4"))
    arch_path, csv_path = workloads.c4_files(d)
    a = ar.Architecture().load_coupling_map(arch_path)
    a.load_properties(csv_path)
    ra = parity.ref().arch_load(arch_path, csv_path)
    return a, ra


def _route(a, circs, kw, options=None, lean=False):
    lib = _abi.load()
    cfg = routing.make_config(debug=True, **kw)
    items = [parity.item_for(lib, a.n_qubits, c, cfg) for c in circs]
    be = routing.CudaBackend(lib, a.handle)
    outs = routing.route_batch(be, a.n_qubits, cfg, items, want_layer_qubits=not lean,
                               options=options, lean=lean)
    return outs, be.stats


def _ref(ra, circ, kw):
    c, t = circ.gate_arrays()
    return parity.ref().map_gates(ra, circ.n_qubits, c, t, parity.refapi.make_config(debug=True, **kw))


def test_config4_heavy_hex_127_fidelity(c4):
    a, ra = c4
    circs = [workloads.c4_circuit(s, 36) for s in range(6)]
    outs, _ = _route(a, circs, workloads.C4_CONFIG)
    big = 0
    for circ, out in zip(circs, outs):
        rr = _ref(ra, circ, workloads.C4_CONFIG)
        assert parity.compare(out, rr, rel=1e-9) == ""
        assert parity.compare(out, rr) == ""  # in fact bit for bit
        assert (out.expanded_nodes, out.generated_nodes) == (rr.expanded_nodes, rr.generated_nodes)
        big = max(big, int(rr.calls["generated_nodes"].max()))
    assert big > 4 * 4096  # open lists long enough for the per-block tops


def test_config4_pool_exhaustion_is_retried(tmp_path, monkeypatch):
    """A first-tier pool far too small for the searches: an explicit
    pool_nodes fails loudly (QMG_POOL_EXHAUSTED); with the automatic pool the
    library routes those instances again with fewer resident searches and
    larger pools and returns what the reference returns."""
    arch_path, csv_path = workloads.c4_files(str(tmp_path))
    ra = parity.ref().arch_load(arch_path, csv_path)
    circs = [workloads.c4_circuit(s, 30) for s in range(40, 48)]
    monkeypatch.setenv("QMAP_B200_POOL_MB", "4")  # 8 searches: ~800 nodes each
    a = ar.Architecture().load_coupling_map(arch_path)  # (fresh workspace)
    a.load_properties(csv_path)
    got, st = _route(a, circs, workloads.C4_CONFIG)
    assert st.retried_instances >= len(circs) // 2
    assert st.kernel_launches >= 2
    assert all(o.status == 0 for o in got)
    for circ, out in list(zip(circs, got))[:3]:
        rr = _ref(ra, circ, workloads.C4_CONFIG)
        assert parity.compare(out, rr) == ""
        assert (out.expanded_nodes, out.generated_nodes) == (rr.expanded_nodes, rr.generated_nodes)
    opts = _abi.QmgRouteOptions()
    opts.pool_nodes = 512
    bad, st = _route(a, circs, workloads.C4_CONFIG, options=opts)
    assert any(o.status == _abi.QMG_POOL_EXHAUSTED for o in bad)
    assert st.retried_instances == 0


def test_config4_batch_is_order_independent(c4):
    a, _ = c4
    circs = [workloads.c4_circuit(s, 24) for s in range(100, 164)]
    outs, _ = _route(a, circs, workloads.C4_CONFIG, lean=True)
    rev, _ = _route(a, circs[::-1], workloads.C4_CONFIG, lean=True)
    for x, y in zip(outs, rev[::-1]):
        assert x.status == 0 and y.status == 0
        assert np.array_equal(x.swaps, y.swaps)
        assert np.array_equal(x.final_qubits, y.final_qubits)
        assert (x.expanded_nodes, x.generated_nodes) == (y.expanded_nodes, y.generated_nodes)


def test_config3_suite_on_heavy_hex_27():
    n, cm = ar.heavy_hex_27()
    a = ar.Architecture(n, cm)
    ra = parity.ref().arch_create(n, sorted(cm))
    inst = workloads.c3_instances(parity.CIRCUITS, 3)
    assert len(inst) >= 60
    circs = [c for _, c in inst]
    outs, _ = _route(a, circs, workloads.C3_CONFIG)
    bad = []
    for (name, circ), out in list(zip(inst, outs))[::2]:
        rr = _ref(ra, circ, workloads.C3_CONFIG)
        msg = parity.compare(out, rr)
        if msg:
            bad.append((name, msg))
    assert not bad, bad[:5]


def _bench_workload(key, **kw):
    import argparse
    import bench
    args = argparse.Namespace(c4_cx=30, cpu_sample=0, searches_per_sm=0)
    for k, v in kw.items():
        setattr(args, k, v)
    return bench, bench.WORKLOADS[key](args)


@pytest.mark.parametrize("key,count", [("c4", 64), ("c3", 400), ("c2", 64)])
def test_bench_workloads_instance_by_instance(key, count, monkeypatch):
    """The very instances bench.py routes (seeds 0..count-1 of each config),
    through the same batched createLayers and qmg_route_batch call, against the
    compiled reference: expanded / generated nodes of the final pass and SWAPs
    emitted per instance, and every routed circuit replayed by the verifier."""
    # (config 4 asks for most of the HBM as node pools; not in a test)
    monkeypatch.setenv("QMAP_B200_POOL_MB", "20000")
    bench, wl = _bench_workload(key)
    L = _abi.load()
    a = wl.architecture(0)
    n = a.n_qubits
    cfg = wl.config()
    seeds = list(range(count))
    cb, items, off, ctrl, tgt = bench.build_instances(L, wl, n, cfg, seeds)
    opts = _abi.QmgRouteOptions()
    wl.options(opts)
    be = routing.CudaBackend(L, a.handle)
    b = routing.pack(n, items, lean=True)
    be.route(cfg, b, opts)
    outs = routing.unpack(b)
    assert all(o.status == 0 for o in outs)
    verdict = cb.verify(a.edges, b.res, len(items), fidelity_aware=cfg.heuristic == 4)
    assert (verdict == 1).all(), np.nonzero(verdict != 1)[0][:10]
    r = bench.run_reference(wl, seeds, min(16, bench.host_cores()))
    assert np.array_equal([o.expanded_nodes_final_pass for o in outs], r["expanded"])
    assert np.array_equal([o.generated_nodes_final_pass for o in outs], r["generated"])
    assert np.array_equal([bench.emitted_swaps(o, cfg) for o in outs], r["swaps"])


def test_verifier_rejects_a_corrupted_routing():
    bench, wl = _bench_workload("c2")
    L = _abi.load()
    a = wl.architecture(0)
    cfg = wl.config()
    cb, items, *_ = bench.build_instances(L, wl, 16, cfg, [0, 1, 2, 3])
    be = routing.CudaBackend(L, a.handle)
    b = routing.pack(16, items, lean=True)
    be.route(cfg, b)
    assert (cb.verify(a.edges, b.res, 4) == 1).all()
    b.swaps[1][3] = (5, 5)                      # a SWAP that is not on a coupling edge
    b.lns[2][5] += 1                            # one SWAP too many in a layer
    v = cb.verify(a.edges, b.res, 4)
    assert v.tolist() == [1, 0, 0, 1]
