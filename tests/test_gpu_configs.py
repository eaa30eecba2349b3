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
