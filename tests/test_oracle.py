"""CPU suite: pins the CPU restatement (oracle/qmap_oracle.cpp) against
 (1) the known-answer vectors of the reference's own tests
     (tests/golden/reference_tests.json, extracted by extract_reference_tests.py),
 (2) outputs of the reference itself committed under tests/golden/
     (make_golden.py ran oracle/_ref), and
 (3) the compiled reference, live, on seeded random circuits when oracle/_ref
     is present (it is built from /root/reference by oracle/Makefile).
"""
import json
import os

import numpy as np
import pytest

import emu_backend
import parity
from oracle import oracleapi
from qmap_b200 import circuit as qcirc
from qmap_b200 import routing

with open(os.path.join(parity.GOLDEN, "reference_tests.json")) as f:
    REFTESTS = json.load(f)

HOSTLIB = None


def hostlib():
    """A library exporting the host-side qmg_circuit_* layering entry points
    (the host debug build; the CUDA library exports the same functions)."""
    global HOSTLIB
    if HOSTLIB is None:
        HOSTLIB = emu_backend.lib()
    return HOSTLIB


_arches = {}


def oracle_arch(name, calib=None):
    key = (name, calib)
    if key not in _arches:
        n, e = parity.read_arch_file(name)
        oa = oracleapi.OracleArch(n, e)
        if calib:
            t = parity.golden_tables()
            oa.set_fidelity_costs(t[f"{name}_c1"], t[f"{name}_c2"], t[f"{name}_cswap"])
        _arches[key] = oa
    return _arches[key]


# ---------------------------------------------------------------------------
# (1) the reference's own known-answer tests
# ---------------------------------------------------------------------------
def test_reference_dijkstra():
    """test/sc/test_general.cpp:63-83"""
    d = REFTESTS["dijkstra"]
    got = oracleapi.build_table(4, d["cm"], d["edge_weights"])
    assert np.array_equal(got, np.array(d["table"], dtype=float))


def test_reference_dijkstra_cnot_reversal():
    """test/sc/test_general.cpp:85-110"""
    d = REFTESTS["dijkstra_cnot_reversal"]
    simple = oracleapi.build_table(5, d["cm"], d["edge_weights"])
    got = oracleapi.build_single_edge_skip_table(simple, d["cm"], d["reversal_cost"])
    assert np.array_equal(got, np.array(d["table"], dtype=float))


def test_reference_dijkstra_skip_edges():
    """test/sc/test_general.cpp:112-192 (five edge-skip levels)"""
    d = REFTESTS["dijkstra_skip_edges"]
    assert np.array_equal(oracleapi.build_table(10, d["cm"], d["edge_weights"]),
                          np.array(d["table"], dtype=float))
    got = oracleapi.build_edge_skip_table(10, d["cm"], d["edge_weights"])
    want = np.array(d["edge_skip_tables"], dtype=float)
    assert got.shape == want.shape
    assert np.array_equal(got, want)


def _route_gates(oa, n_logical, cx, **kw):
    circ = qcirc.Circuit(n_logical, [qcirc.Op("cx", (c, t)) for c, t in cx])
    # the reference's test circuits measure every qubit, so none of them is
    # idle and the identity layout places all of them (Mapper.cpp:44-53)
    circ.measurements = [(q, q) for q in range(n_logical)]
    cfg = routing.make_config(debug=True, **kw)
    item = parity.item_for(hostlib(), oa.n, circ, cfg)
    return routing.route_batch(oracleapi.OracleBackend(oa), oa.n, cfg, [item],
                               want_layer_qubits=True)[0]


def test_reference_node_counts_ring5():
    """test/sc/heuristic/test_heuristic.cpp:776-855"""
    d = REFTESTS["node_counts"]["ring5_three_cx"]
    oa = oracleapi.OracleArch(d["n"], d["cm"])
    out = _route_gates(oa, d["n"], d["cx"], **d["settings"])
    assert out.status == 0 and out.n_layers_done == d["layers"]
    assert list(out.layers["solution_depth"]) == d["solution_depth"]
    assert list(out.layers["generated_nodes"]) == d["generated_nodes"]
    assert list(out.layers["expanded_nodes"]) == d["expanded_nodes"]
    assert out.generated_nodes == d["total_generated"]
    assert out.expanded_nodes == d["total_expanded"]


def test_reference_generated_nodes_qx5():
    """test/sc/heuristic/test_heuristic.cpp:869-893"""
    d = REFTESTS["node_counts"]["qx5_single_cx"]
    out = _route_gates(oracle_arch(d["arch"]), d["n"], d["cx"], **d["settings"])
    assert list(out.layers["generated_nodes"]) == d["generated_nodes"]
    assert out.generated_nodes == d["total_generated"]


def _trim(q):
    """test_heuristic.cpp:705-716: the layout up to its last occupied entry"""
    last = 0
    for j in range(len(q) - 1, 0, -1):
        if q[j] != -1:
            last = j
            break
    return [int(v) for v in q[: last + 1]]


OPT = REFTESTS["optimal_solutions"]


@pytest.mark.parametrize("heuristic", OPT["heuristics"])
@pytest.mark.parametrize("circuit", sorted(OPT["layouts"]))
def test_reference_optimal_solutions(circuit, heuristic):
    """test/sc/heuristic/test_heuristic.cpp:276-556 checked as at :699-721: the
    tight, principally admissible heuristics reproduce the listed per-layer
    layouts on ibmq_yorktown, ibmq_london and ibm_qx5 (in that order)."""
    circ = qcirc.load_qasm(os.path.join(parity.CIRCUITS, circuit + ".qasm"))
    cfg = routing.make_config(debug=True, heuristic=heuristic, **OPT["settings"])
    got = []
    for arch in OPT["architectures_in_order"]:
        oa = oracle_arch(arch)
        if circ.n_qubits > oa.n:
            continue
        item = parity.item_for(hostlib(), oa.n, circ, cfg)
        out = routing.route_batch(oracleapi.OracleBackend(oa), oa.n, cfg, [item],
                                  want_layer_qubits=True)[0]
        assert out.status == 0
        got += [_trim(q) for q in out.layer_qubits]
    assert got == OPT["layouts"][circuit]


# ---------------------------------------------------------------------------
# (2) committed outputs of the reference
# ---------------------------------------------------------------------------
def test_tables_match_reference_outputs():
    t = parity.golden_tables()
    names = sorted(k[:-5] for k in t.files if k.endswith("_dist"))
    assert len(names) >= 8
    for name in names:
        oa = oracle_arch(name)
        assert np.array_equal(oa.distance_table(False), t[f"{name}_dist"]), name
        assert np.array_equal(oa.distance_table(True), t[f"{name}_dist_rev"]), name
    oa = oracle_arch("ibmq_london", "ibmq_london.csv")
    assert np.array_equal(oa.fidelity_distance_tables(), t["ibmq_london_fid"])


SLOW_NODES = 50000  # cases above this run with QMAP_SLOW=1 (minutes on CPU)
RUNS = parity.golden_runs()
FAST = [g for g in RUNS if g.generated_nodes <= SLOW_NODES]
SLOW = [g for g in RUNS if g.generated_nodes > SLOW_NODES]


def _check_golden(chunk):
    bad = []
    for g in chunk:
        oa = oracle_arch(g.meta["arch"], g.meta["calibration"])
        msg, _ = parity.run_golden(oracleapi.OracleBackend(oa), hostlib(), oa.n, g)
        if msg:
            bad.append((g.label, msg))
    assert not bad, bad[:5]


@pytest.mark.parametrize("part", range(8))
def test_route_matches_reference_outputs(part):
    """every aStarMap call of the final pass: 4 costs, 7 counters, result node
    id, result layout, SWAP list (bit-exact, fidelity costs included)"""
    assert len(FAST) > 250
    _check_golden(FAST[part::8])


@pytest.mark.skipif(not os.environ.get("QMAP_SLOW"), reason="minutes of CPU; set QMAP_SLOW=1")
def test_route_matches_reference_outputs_slow():
    _check_golden(SLOW)


def test_split_cases_present():
    assert sum(g.meta["split"] for g in FAST) >= 20


# ---------------------------------------------------------------------------
# (3) live against the compiled reference
# ---------------------------------------------------------------------------
@pytest.mark.skipif(not parity.have_ref(), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("arch", ["ibm_qx5", "ibmq_tokyo"])
def test_random_circuits_against_compiled_reference(arch):
    ra = parity.ref().arch_load(os.path.join(parity.ARCHS, arch + ".arch"))
    oa = oracle_arch(arch)
    be = oracleapi.OracleBackend(oa)
    bad = []
    for seed in range(12):
        # (individual gates keep every search small; the disjoint layerings of
        # random circuits can take the reference minutes per circuit)
        circ = qcirc.random_circuit(ra.n - (seed % 3), 30 + 3 * seed, seed=seed)
        kw = dict(heuristic=seed % 4, layering="individual_gates",
                  initial_layout=("dynamic", "identity")[seed % 2],
                  lookahead_heuristic=("gate_count_max_distance", "none",
                                       "gate_count_sum_distance")[seed % 3],
                  iterative_bidirectional_routing_passes=seed % 2)
        msg, _, rr = parity.run_pair(be, hostlib(), ra, circ, **kw)
        if msg and "split" not in msg:
            bad.append((seed, kw, msg))
    assert not bad, bad[:3]


# ---------------------------------------------------------------------------
# the reference's own Architecture tests (test/sc/test_architecture.cpp:274-960)
# ---------------------------------------------------------------------------
def _arch_goldens():
    import json
    with open(os.path.join(parity.ROOT, "tests", "golden", "architecture_tests.json")) as f:
        return json.load(f)["tests"]


def check_architecture_golden(g, tables, swap_cost):
    """tables: (K, n, n) k-edge-skip fidelity distance tables, swap_cost (n, n);
    a table index >= K reads as all zero (Architecture.hpp:312-314)."""
    n = g["n"]
    for t in g["fidelity_distance_tables"]:
        got = tables[t["skip"]] if t["skip"] < len(tables) else np.zeros((n, n))
        np.testing.assert_allclose(got, np.array(t["table"]), rtol=0, atol=t["tol"],
                                   err_msg=f"{g['test']} skip {t['skip']}")
    src = {"swapFidCost": swap_cost, "fidDistance": tables[0]}
    for e in g["near"]:
        assert abs(src[e["what"]][e["i"]][e["j"]] - e["value"]) <= e["tol"], (g["test"], e)
    for e in g["greater_than_1e20"]:
        assert src[e["what"]][e["i"]][e["j"]] > 1e20, (g["test"], e)


@pytest.mark.parametrize("g", _arch_goldens(), ids=lambda g: g["test"])
def test_compiled_reference_reproduces_its_architecture_tests(g):
    ra = parity.ref().arch_create(g["n"], [tuple(e) for e in g["coupling_map"]])
    rate = {(a, b): r for a, b, r in g["two_qubit_error"]}
    ra.set_error_rates(g["single_qubit_error"], [rate[tuple(int(x) for x in e)] for e in ra.edges])
    check_architecture_golden(g, ra.fidelity_distance_tables(), ra.swap_fidelity_costs())


def test_distance_cheapest_path_reference():
    """test_architecture.cpp:928-960 (DistanceCheapestPathTest): 9 unidirectional
    edges (34 each) lose against 10 bidirectional ones (30 each)."""
    n, cm, nr = parity.cheapest_path_architecture()
    ra = parity.ref().arch_create(n, cm)
    d = ra.distance_table(True)
    assert d.shape == (2 * nr + 1, 2 * nr + 1)
    assert abs(d[0][nr] - nr * 30.0) <= 1e-6
