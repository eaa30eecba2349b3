"""-m gpu: the CUDA library (through the C ABI) against the committed outputs
of the reference (tests/golden/route_golden.npz, tables_golden.npz) - these
tests need no oracle library on the GPU box."""
import os

import numpy as np
import pytest

import parity
from qmap_b200 import _abi, routing
from qmap_b200 import architecture as ar
from qmap_b200 import circuit as qcirc

pytestmark = pytest.mark.gpu

_archs = {}


def arch_for(name, calib=None):
    key = (name, calib)
    if key not in _archs:
        a = ar.Architecture().load_coupling_map(os.path.join(parity.ARCHS, name + ".arch"))
        if calib:
            a.load_properties(os.path.join(parity.ARCHS, calib))
        _archs[key] = a
    return _archs[key]


def test_tables_match_reference_outputs():
    t = parity.golden_tables()
    names = sorted(k[:-5] for k in t.files if k.endswith("_dist"))
    for name in names:
        a = arch_for(name)
        assert np.array_equal(a.distance_table(False), t[f"{name}_dist"]), name
        assert np.array_equal(a.distance_table(True), t[f"{name}_dist_rev"]), name
    a = arch_for("ibmq_london", "ibmq_london.csv")
    assert np.array_equal(a.fidelity_distance_tables(), t["ibmq_london_fid"])


RUNS = [g for g in parity.golden_runs() if not g.meta["split"]]


@pytest.mark.parametrize("part", range(8))
def test_route_matches_reference_outputs(part):
    """every final-pass aStarMap call: costs, counters, result node id, layout,
    SWAP list - bit-exact, fidelity costs included"""
    assert len(RUNS) > 270
    L = _abi.load()
    bad = []
    for g in RUNS[part::8]:
        a = arch_for(g.meta["arch"], g.meta["calibration"])
        be = routing.CudaBackend(L, a.handle)
        kw = dict(g.meta["config"], debug=True)
        cfg = routing.make_config(**kw)
        circ = qcirc.load_qasm(os.path.join(parity.CIRCUITS, g.meta["circuit"]))
        opts = None
        if g.generated_nodes > 100000:
            opts = _abi.QmgRouteOptions()
            opts.pool_nodes = 1 << 20
        out = routing.route_batch(be, a.n_qubits, cfg, [parity.item_for(L, a.n_qubits, circ, cfg)],
                                  want_layer_qubits=True, options=opts)[0]
        msg = parity.compare(out, g)
        if msg:
            bad.append((g.label, msg))
    assert not bad, bad[:5]


def test_batch_of_all_cases_per_configuration():
    """many instances in ONE launch: all golden cases that share an architecture
    and a configuration are routed as one batch (lean outputs) and every
    instance's SWAP list / final layout / counters are checked"""
    L = _abi.load()
    groups = {}
    for g in RUNS:
        if g.generated_nodes > 100000:
            continue
        key = (g.meta["arch"], g.meta["calibration"], tuple(sorted(g.meta["config"].items())))
        groups.setdefault(key, []).append(g)
    checked = 0
    for (arch, calib, kw), gs in groups.items():
        a = arch_for(arch, calib)
        cfg = routing.make_config(**dict(kw))
        items = []
        for g in gs:
            circ = qcirc.load_qasm(os.path.join(parity.CIRCUITS, g.meta["circuit"]))
            items.append(parity.item_for(L, a.n_qubits, circ, cfg))
        # the same instances several times over, so that warps take >1 instance
        reps = 3
        outs = routing.route_batch(routing.CudaBackend(L, a.handle), a.n_qubits, cfg,
                                   items * reps, lean=True)
        for i, o in enumerate(outs):
            g = gs[i % len(gs)]
            assert o.status == 0, g.label
            assert o.layers is None
            assert np.array_equal(o.layer_n_swaps, g.calls["n_swaps"]), g.label
            assert np.array_equal(o.swaps, g.swap_pairs), g.label
            if g.n_layers:
                assert np.array_equal(o.final_qubits, g.qubits_after[-1]), g.label
            checked += 1
    assert checked > 600


def test_swap_buffer_too_small_is_reported_and_retried():
    L = _abi.load()
    g = next(g for g in RUNS if g.meta["circuit"] == "qft_16.qasm" and g.meta["arch"] == "ibm_qx5"
             and not g.meta["config"])
    a = arch_for("ibm_qx5")
    cfg = routing.make_config()
    circ = qcirc.load_qasm(os.path.join(parity.CIRCUITS, "qft_16.qasm"))
    item = parity.item_for(L, a.n_qubits, circ, cfg)
    be = routing.CudaBackend(L, a.handle)
    b = routing.pack(a.n_qubits, [item], swap_capacity=8, lean=True)
    be.route(cfg, b)
    out = routing.unpack(b)[0]
    assert out.status == _abi.QMG_BUFFER_TOO_SMALL
    assert int(b.res[0].n_swaps_total) > 8  # the search stopped where the buffer overflowed
    full = routing.route_batch(be, a.n_qubits, cfg, [item], lean=True)[0]
    assert full.status == 0 and len(full.swaps) == g.swaps


def test_workspace_is_reused_between_batches():
    """the second call on one Architecture reuses the staging buffers and gives
    identical results; a bigger batch afterwards grows them"""
    L = _abi.load()
    a = arch_for("ibm_qx5")
    cfg = routing.make_config(initial_layout="identity")
    circ = qcirc.load_qasm(os.path.join(parity.CIRCUITS, "qft_16.qasm"))
    item = parity.item_for(L, a.n_qubits, circ, cfg)
    be = routing.CudaBackend(L, a.handle)
    first = routing.route_batch(be, a.n_qubits, cfg, [item] * 8, lean=True)
    second = routing.route_batch(be, a.n_qubits, cfg, [item] * 8, lean=True)
    big = routing.route_batch(be, a.n_qubits, cfg, [item] * 700, lean=True)
    for o in second + big:
        assert o.status == 0
        assert np.array_equal(o.swaps, first[0].swaps)
        assert o.expanded_nodes == first[0].expanded_nodes


def test_mapper_mirror_on_cuda():
    """HeuristicMapper.map() end to end on the CUDA library: routed circuits
    verify against the input, SWAP counts equal the reference's, a low split
    limit exercises split-and-resume."""
    from qmap_b200 import mapper
    a = arch_for("ibm_qx5")
    checked = 0
    for g in RUNS:
        if g.meta["arch"] != "ibm_qx5" or g.generated_nodes > 20000:
            continue
        circ = qcirc.load_qasm(os.path.join(parity.CIRCUITS, g.meta["circuit"]))
        m = mapper.HeuristicMapper(circ, a)
        res = m.map(**g.meta["config"])
        assert mapper.verify_mapping(circ, m) == "", g.label
        assert res.output_swaps == g.swaps, g.label
        checked += 1
    assert checked > 40
    circ = qcirc.load_qasm(os.path.join(parity.CIRCUITS, "qft_16.qasm"))
    m = mapper.HeuristicMapper(circ, a)
    res = m.map(layering="disjoint_qubits", automatic_layer_splits_node_limit=3)
    assert res.layer_splits > 0
    assert mapper.verify_mapping(circ, m) == ""


class _RawLayers:
    """qmg_layers built by hand (what a C caller may pass): not necessarily the
    product of createLayers."""

    def __init__(self, n_logical, layers2, layers1):
        import ctypes as C
        self.off2 = np.zeros(len(layers2) + 1, dtype=np.uint32)
        self.off1 = np.zeros(len(layers2) + 1, dtype=np.uint32)
        g2, g1 = [], []
        for l, (a, b) in enumerate(zip(layers2, layers1)):
            g2 += a
            g1 += b
            self.off2[l + 1] = len(g2)
            self.off1[l + 1] = len(g1)
        self.g2 = np.array(g2, dtype=np.uint16).reshape(-1, 4)
        self.g1 = np.array(g1, dtype=np.uint16).reshape(-1, 2)
        v = _abi.QmgLayers()
        v.n_layers = len(layers2)
        v.n_logical = n_logical
        v.gate2q_offsets = self.off2.ctypes.data
        v.gates2q = self.g2.ctypes.data if len(self.g2) else None
        v.gate1q_offsets = self.off1.ctypes.data
        v.gates1q = self.g1.ctypes.data if len(self.g1) else None
        self.layers = v
        self.n_layers = len(layers2)


def test_hand_made_layers_ragged_and_empty():
    """layers a C caller may hand over: empty layers, layers whose gates share a
    qubit, only-1q layers, an instance with no layers at all - the tables the
    device derives from them give the oracle's results"""
    from oracle import oracleapi
    L = _abi.load()
    a = arch_for("ibmq_tokyo")
    n = a.n_qubits
    oa = oracleapi.OracleArch(n, a.edges)
    raw = _RawLayers(
        8,
        [[], [[0, 5, 1, 0], [5, 7, 0, 2]], [], [[1, 6, 1, 1]], [[0, 7, 2, 0], [2, 3, 1, 0], [4, 6, 0, 1]],
         [], [[3, 5, 1, 0]]],
        [[[2, 1]], [], [], [[1, 2], [3, 1], [4, 1]], [], [[0, 3]], []])
    empty = _RawLayers(4, [], [])
    for kw in (dict(), dict(heuristic=2, lookahead_heuristic="gate_count_sum_distance"),
               dict(heuristic=3, initial_layout="identity", automatic_layer_splits=False)):
        cfg = routing.make_config(debug=True, **kw)
        init = np.arange(n, dtype=np.int16) if cfg.initial_layout == 0 else None
        items = [routing.RouteItem(raw, init), routing.RouteItem(empty, init),
                 routing.RouteItem(raw, init, 1)]
        got = routing.route_batch(routing.CudaBackend(L, a.handle), n, cfg, items,
                                  want_layer_qubits=True)
        want = routing.route_batch(oracleapi.OracleBackend(oa), n, cfg, items,
                                   want_layer_qubits=True)
        for g, w in zip(got, want):
            assert parity.compare_outputs(g, w) == "", kw
        assert got[1].n_layers_done == 0 and got[1].status == 0


SPLIT_RUNS = [g for g in parity.golden_runs() if g.meta["split"]]


def test_layer_splits_match_reference_outputs():
    """every case in which the reference split a layer (final pass and pseudo-
    routing passes): the kernel splits the layer itself (Mapper::splitLayer on
    its private copy of the circuit tables) and restarts on the uncleared open
    list - all calls agree bit for bit, the split log replays to the
    reference's layer structure"""
    assert len(SPLIT_RUNS) >= 30
    L = _abi.load()
    bad, splits = [], 0
    for g in SPLIT_RUNS:
        a = arch_for(g.meta["arch"], g.meta["calibration"])
        be = routing.CudaBackend(L, a.handle)
        opts = None
        if g.generated_nodes > 100000:
            opts = _abi.QmgRouteOptions()
            opts.pool_nodes = 1 << 20
        msg, out = parity.run_golden(be, L, a.n_qubits, g, options=opts)
        if msg:
            bad.append((g.label, msg))
            continue
        splits += len(out.split_layers)
        cfg = routing.make_config(**g.meta["config"])
        circ = qcirc.load_qasm(os.path.join(parity.CIRCUITS, g.meta["circuit"]))
        lc = parity.item_for(L, a.n_qubits, circ, cfg).circuit
        if len(out.split_layers):
            lc = lc.split_copy(out.split_layers)
        assert lc.n_layers == g.n_layers, g.label
    assert not bad, bad[:3]
    assert splits > 100


def test_batch_with_layer_splits_against_oracle():
    """one launch, many instances that split layers (seeded random circuits,
    tiny split limits, pseudo-routing passes): lean outputs and split logs equal
    the CPU restatement's, instance by instance"""
    from oracle import oracleapi
    L = _abi.load()
    for arch, calib, heurs in (("ibm_qx5", None, (0, 1, 2, 3)), ("ibmq_tokyo", None, (0, 3)),
                               ("ibmq_london", "ibmq_london.csv", (4,))):
        a = arch_for(arch, calib)
        n = a.n_qubits
        oa = oracleapi.OracleArch(n, a.edges)
        if calib:
            t = parity.golden_tables()
            oa.set_fidelity_costs(t[f"{arch}_c1"], t[f"{arch}_c2"], t[f"{arch}_cswap"])
        for heur in heurs:
            for passes in (0, 1):
                cfg = routing.make_config(
                    heuristic=heur, layering=("disjoint_qubits", "disjoint_2q_blocks")[heur % 2],
                    lookahead_heuristic="none" if heur == 4 else "gate_count_max_distance",
                    iterative_bidirectional_routing_passes=passes,
                    automatic_layer_splits_node_limit=(2, 3, 5)[heur % 3])
                items = []
                for seed in range(24):
                    circ = qcirc.random_circuit(n - (seed % 3) if n > 6 else n, 24 + 2 * seed,
                                                seed=1000 * heur + seed)
                    items.append(parity.item_for(L, n, circ, cfg))
                got = routing.route_batch(routing.CudaBackend(L, a.handle), n, cfg, items * 3,
                                          lean=True, extra_layers=160)
                want = routing.route_batch(oracleapi.OracleBackend(oa), n, cfg, items,
                                           lean=True, extra_layers=160)
                nsplit = 0
                for i, o in enumerate(got):
                    w = want[i % len(items)]
                    key = (arch, heur, passes, i)
                    assert o.status == 0 and o.n_layers_done == w.n_layers_done, key
                    assert np.array_equal(o.split_layers, w.split_layers), key
                    assert np.array_equal(o.layer_n_swaps, w.layer_n_swaps), key
                    assert np.array_equal(o.swaps, w.swaps), key
                    assert np.array_equal(o.final_qubits, w.final_qubits), key
                    assert (o.expanded_nodes, o.generated_nodes) == \
                        (w.expanded_nodes, w.generated_nodes), key
                    nsplit += len(w.split_layers)
                assert nsplit > 0, (arch, heur, passes)


@pytest.mark.parametrize("arch,calib,heuristics,circuits", [
    ("ibmq_yorktown.arch", None, [0, 1, 2, 3], parity.DATALOG_CIRCUITS),
    ("ibmq_london.arch", "ibmq_london.csv", [0, 4], parity.DATALOG_CIRCUITS),
    ("ibm_qx5.arch", None, [0, 1, 2, 3], parity.DATALOG_CIRCUITS + ["ising_model_10", "rd73_140"]),
])
def test_node_log_equals_the_reference_datalogger(arch, calib, heuristics, circuits):
    """Row f3 on the device: every node generated in the HeuristicProperties
    fixture runs (test/sc/heuristic/test_heuristic.cpp:224-775; 9 circuits,
    IBM Yorktown / London with calibration / QX5), written in the DataLogger's
    CSV schema (src/sc/DataLogger.cpp:216-269), equals the files the compiled
    reference writes; the fixture's property checks run on the device's log."""
    from qmap_b200 import architecture as ar2
    a = ar2.Architecture().load_coupling_map(os.path.join(parity.ARCHS, arch))
    if calib:
        a.load_properties(os.path.join(parity.ARCHS, calib))
    lib = _abi.load()
    be = routing.CudaBackend(lib, a.handle)
    total = 0
    for circuit in circuits:
        if qcirc.load_qasm(os.path.join(parity.CIRCUITS, circuit + ".qasm")).n_qubits > a.n_qubits:
            continue
        for h in heuristics:
            total += parity.check_node_log(be, lib, arch, calib, circuit, h)
    assert total > 300
