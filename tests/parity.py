"""Shared parity helpers: run one routing job through a backend (CUDA library
or the host debug build) and through the oracle (the compiled reference,
oracle/_ref) and compare every aStarMap call: costs, node counters, result
node id, result layout and SWAP list."""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "mqt-qmap_b200"), os.path.join(ROOT, "tests", "emu")):
    if p not in sys.path:
        sys.path.insert(0, p)

from oracle import refapi  # noqa: E402
from qmap_b200 import circuit as qcirc  # noqa: E402
from qmap_b200 import routing  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
CIRCUITS = os.path.join(GOLDEN, "circuits")
ARCHS = os.path.join(GOLDEN, "arch")

KEYS = ("cost_fixed", "cost_fixed_reversals", "cost_heur", "lookahead_penalty",
        "expanded_nodes", "generated_nodes", "solution_depth", "result_node_id",
        "expanded_nodes_after_first_solution", "expanded_nodes_after_optimal_solution",
        "solution_nodes", "solution_nodes_after_optimal_solution", "early_termination")

_ref = None


DBL_MAX = np.finfo(np.float64).max


def random_weighted(n, extra, seed, integer):
    """Connected random coupling map on n qubits (a spanning tree + `extra`
    edges, each with probability 1/2 in both directions) and random directed
    weights; returns (sorted coupling entries, n x n weights, DBL_MAX off-edge)."""
    rng = np.random.default_rng(seed)
    und = {(i, int(rng.integers(0, i))) for i in range(1, n)}
    while len(und) < n - 1 + extra:
        a, b = (int(x) for x in rng.integers(0, n, 2))
        if a != b and (b, a) not in und:
            und.add((a, b))
    w = np.full((n, n), DBL_MAX)
    cm = []
    for a, b in und:
        cm.append((a, b))
        if rng.random() < 0.5:
            cm.append((b, a))
        for x, y in ((a, b), (b, a)):
            w[x, y] = float(rng.integers(1, 40)) if integer else float(rng.uniform(0.01, 3.0))
    return sorted(set(cm)), w


def cheapest_path_architecture():
    """The architecture of test_architecture.cpp:928-960: a path of nrEdges
    unidirectional edges 0..nrEdges and a detour of nrEdges + 1 bidirectional
    ones; nrEdges = 1 + ceil(30 / (34 - 30)) = 9."""
    import math
    nr = 1 + math.ceil(30 / (34 - 30))
    cm = [(i + 1, i) for i in range(nr)]
    for i in range(nr + 1, 2 * nr):
        cm += [(i, i + 1), (i + 1, i)]
    cm += [(0, nr + 1), (nr + 1, 0), (2 * nr, nr), (nr, 2 * nr)]
    return 2 * nr + 1, sorted(set(cm)), nr


def ref() -> refapi.Ref:
    global _ref
    if _ref is None:
        _ref = refapi.Ref()
    return _ref


_ref_archs = {}


def ref_arch_cached(n, cm):
    """parity.ref().arch_create, kept per process (the reference needs about a
    minute for the tables of a 32 x 32 grid)."""
    key = (n, tuple(sorted(cm)))
    if key not in _ref_archs:
        _ref_archs[key] = ref().arch_create(n, sorted(cm))
    return _ref_archs[key]


def have_ref() -> bool:
    return os.path.exists(refapi.REF_LIB)


def item_for(lib, n, circ, cfg):
    c, t, identity = circ.mapper_view(n)
    lc = routing.LayeredCircuit(lib, n, c, t, cfg.layering)
    init = identity if cfg.initial_layout == 0 else None
    return routing.RouteItem(lc, init, cfg.iterative_bidirectional_routing_passes)


def compare(out, rr, rel=0.0, allow_split=False):
    """out: routing.RouteOutput, rr: refapi.RefMapResult.  Returns '' if every
    final-pass aStarMap call agrees, else a description of the first
    difference.  rel > 0 allows that relative error on the four cost fields."""
    nl = rr.n_layers
    if (rr.call_flags & 2).any() and not allow_split:
        return "reference split a layer (not comparable call by call)"
    if out.status != 0:
        return f"status {out.status}"
    if out.n_layers_done != nl:
        return f"layers done {out.n_layers_done} != {nl}"
    # a call that split its layer is recorded twice by the reference driver
    # (the restarted inner search, then the outer call flagged "split")
    keep = np.nonzero((rr.call_flags & 2) == 0)[0][-nl:] if nl else np.zeros(0, dtype=np.int64)
    calls = rr.calls[keep]
    qa = rr.qubits_after[keep]
    debug = bool(calls["generated_nodes"].any())
    for l in range(nl):
        a, b = out.layers[l], calls[l]
        for k in KEYS:
            if not debug and k not in KEYS[:4] + ("solution_depth", "result_node_id"):
                continue
            av, bv = a[k], b[k]
            if av == bv:
                continue
            if k in KEYS[:4] and rel > 0 and abs(av - bv) <= rel * max(abs(av), abs(bv)):
                continue
            return f"layer {l}: {k} {av} != {bv}"
        if out.layer_qubits is not None and not (out.layer_qubits[l] == qa[l]).all():
            return f"layer {l}: layout {out.layer_qubits[l]} != {qa[l]}"
        rs = rr.swap_pairs[b["swap_offset"]: b["swap_offset"] + b["n_swaps"]]
        if not np.array_equal(out.layer_swaps(l), rs):
            return f"layer {l}: swaps differ"
    # (rr.final_qubits is not comparable: routeCircuit rewinds `qubits` through
    # the SWAPs when it fixes up early single-qubit gates, :461-508)
    if nl > 0 and not np.array_equal(out.final_qubits, qa[nl - 1]):
        return "final layout differs"
    return ""


def run_pair(backend, lib, ra, circ_path_or_circ, rel=0.0, **kw):
    """Route with `backend` and with the reference; returns (msg, out, rr)."""
    kw.setdefault("debug", True)
    cfg_r = refapi.make_config(**kw)
    cfg = routing.make_config(**kw)
    if isinstance(circ_path_or_circ, str):
        circ = qcirc.load_qasm(circ_path_or_circ)
        rr = ref().map_qasm(ra, circ_path_or_circ, cfg_r)
    else:
        circ = circ_path_or_circ
        if circ.measurements:
            # (measured qubits are not idle: the reference must see them)
            import tempfile
            with tempfile.NamedTemporaryFile("w", suffix=".qasm", delete=False) as f:
                f.write(circ.to_qasm())
            try:
                rr = ref().map_qasm(ra, f.name, cfg_r)
            finally:
                os.unlink(f.name)
        else:
            c, t = circ.gate_arrays()
            rr = ref().map_gates(ra, circ.n_qubits, c, t, cfg_r)
    # room for the layers a split adds: the backend then splits layers itself
    # exactly where the reference does (qmg_instance.layer_capacity)
    split = bool((rr.call_flags & 2).any())
    out = routing.route_batch(backend, ra.n, cfg, [item_for(lib, ra.n, circ, cfg)],
                              want_layer_qubits=True,
                              extra_layers=rr.n_layers if split else 0)[0]
    return compare(out, rr, rel, allow_split=True), out, rr


def cuda_backend_for(ra, arch=None):
    """CudaBackend on a qmap_b200 Architecture equal to the reference arch."""
    from qmap_b200 import _abi
    from qmap_b200.architecture import Architecture
    if arch is None:
        arch = Architecture(ra.n, [tuple(e) for e in ra.edges])
    return routing.CudaBackend(_abi.load(), arch.handle), arch


def compare_outputs(a, b) -> str:
    """Two routing.RouteOutput objects (e.g. CUDA vs the CPU restatement)."""
    if a.status != b.status or a.n_layers_done != b.n_layers_done:
        return f"status/layers {a.status}/{a.n_layers_done} != {b.status}/{b.n_layers_done}"
    for l in range(a.n_layers_done):
        for k in KEYS + ("n_swaps",):
            if a.layers[l][k] != b.layers[l][k]:
                return f"layer {l}: {k} {a.layers[l][k]} != {b.layers[l][k]}"
        if not np.array_equal(a.layer_swaps(l), b.layer_swaps(l)):
            return f"layer {l}: swaps differ"
        if a.layer_qubits is not None and b.layer_qubits is not None and \
                not np.array_equal(a.layer_qubits[l], b.layer_qubits[l]):
            return f"layer {l}: layout differs"
    if not np.array_equal(a.final_qubits, b.final_qubits):
        return "final layout differs"
    return ""


def smoke_check() -> int:
    """Used by __graft_entry__.smoke(): configs 1 and 2 on cuda:0 against the
    oracle (the CPU restatement, itself pinned to the reference) and against
    the committed outputs of the reference."""
    from oracle import oracleapi
    from qmap_b200 import _abi
    from qmap_b200.architecture import Architecture
    lib = _abi.load()
    total = 0
    for arch_name, qasm, kw in (
            ("ibm_qx4", "4gt13_92.qasm", dict(lookahead_heuristic="none")),
            ("ibm_qx5", "qft_16.qasm", dict())):
        n, edges = read_arch_file(arch_name)
        arch = Architecture(n, [tuple(int(v) for v in e) for e in edges])
        oa = oracleapi.OracleArch(n, edges)
        for inc in (False, True):
            if not np.array_equal(arch.distance_table(inc), oa.distance_table(inc)):
                raise AssertionError(f"{arch_name}: distance table differs from the oracle")
        cfg = routing.make_config(debug=True, **kw)
        circ = qcirc.load_qasm(os.path.join(CIRCUITS, qasm))
        item = item_for(lib, n, circ, cfg)
        out = routing.route_batch(routing.CudaBackend(lib, arch.handle), n, cfg, [item],
                                  want_layer_qubits=True)[0]
        want = routing.route_batch(oracleapi.OracleBackend(oa), n, cfg, [item],
                                   want_layer_qubits=True)[0]
        msg = compare_outputs(out, want)
        if msg:
            raise AssertionError(f"{qasm} on {arch_name} vs the oracle: {msg}")
        g = next(g for g in golden_runs() if g.meta["arch"] == arch_name
                 and g.meta["circuit"] == qasm and g.meta["config"] == kw)
        msg = compare(out, g)
        if msg:
            raise AssertionError(f"{qasm} on {arch_name} vs the reference's output: {msg}")
        total += out.n_layers_done
    return total


# ---------------------------------------------------------------------------
# committed golden vectors (tests/golden/make_golden.py ran the reference)
# ---------------------------------------------------------------------------
class GoldenRun:
    """The slice of refapi.RefMapResult `compare` reads, from the fixtures."""

    def __init__(self, meta, calls, qubits, swaps):
        self.meta = meta
        self.n_layers = int(meta["layers"])
        self.calls = calls
        self.qubits_after = qubits
        self.swap_pairs = swaps
        self.call_flags = np.zeros(len(calls), dtype=np.int32)
        self.expanded_nodes = int(meta["expanded_nodes"])
        self.generated_nodes = int(meta["generated_nodes"])
        self.swaps = int(meta["swaps"])

    @property
    def label(self):
        m = self.meta
        kw = ",".join(f"{k}={v}" for k, v in sorted(m["config"].items()))
        return f"{m['arch']}/{m['circuit']}[{kw}]"


_golden = None


def golden_runs():
    """All committed golden routing runs (GoldenRun), in manifest order."""
    global _golden
    if _golden is None:
        import json
        with open(os.path.join(GOLDEN, "route_golden.json")) as f:
            manifest = json.load(f)
        z = np.load(os.path.join(GOLDEN, "route_golden.npz"))
        _golden = [GoldenRun(m, z[f"c{i}_calls"], z[f"c{i}_qubits"], z[f"c{i}_swaps"])
                   for i, m in enumerate(manifest)]
    return _golden


def golden_tables():
    return np.load(os.path.join(GOLDEN, "tables_golden.npz"))


def read_arch_file(name):
    """(n, edges) of a tests/golden/arch/<name>.arch fixture."""
    with open(os.path.join(ARCHS, name + ".arch")) as f:
        tok = f.read().split()
    n = int(tok[0])
    e = np.array(tok[1:], dtype=np.int64).reshape(-1, 2)
    return n, e.astype(np.uint16)


def run_golden(backend, lib, n, g: GoldenRun, rel=0.0, backend_splits=True, options=None):
    """Route the golden case `g` with `backend`; returns (msg, out).  Cases in
    which the reference split a layer get room for the split layers, which
    makes the backend split layers itself (qmg_instance.layer_capacity)."""
    kw = dict(g.meta["config"])
    kw["debug"] = True
    cfg = routing.make_config(**kw)
    circ = qcirc.load_qasm(os.path.join(CIRCUITS, g.meta["circuit"]))
    out = routing.route_batch(backend, n, cfg, [item_for(lib, n, circ, cfg)],
                              want_layer_qubits=True, options=options,
                              extra_layers=g.n_layers if g.meta["split"] and backend_splits
                              else 0)[0]
    return compare(out, g, rel, allow_split=True), out


def fidelity_known_answers():
    """The reference's fidelity-aware known-answer tests, re-expressed
    (test/sc/heuristic/test_heuristic.cpp:1769-1851 RemapSingleQubit, :1853-1932
    QubitRideAlong, :1934-1981 SingleQubitsCompete): architecture, error rates,
    circuit (every qubit measured, so none is idle), expected SWAP count and
    total log-fidelity cost.  One layer each (disjoint_2q_blocks), identity
    layout, swapOnFirstLayer."""
    import math

    def c(e):
        return -math.log2(1 - e)

    def path(n):
        return [(i, i + 1) for i in range(n - 1)] + [(i + 1, i) for i in range(n - 1)]

    def sym(d):
        return {**d, **{(b, a): r for (a, b), r in d.items()}}

    e5, e4, e3, e1, e0 = 0.99, 0.9, 0.5, 0.1, 0.01
    out = []
    circ = qcirc.Circuit(6, measurements=[(i, i) for i in range(6)], n_clbits=6)
    for _ in range(5):
        circ.cx(2, 0)
        circ.gate("x", 3)
    out.append(dict(name="RemapSingleQubit", n=6, cm=path(6), q1=[e5, e5, e5, e4, e4, e1],
                    cx=sym({(0, 1): e1, (1, 2): e3, (2, 3): e5, (3, 4): e0, (4, 5): e0}), circuit=circ,
                    swaps=3, fidelity=3 * c(e3) + 3 * c(e0) + 3 * c(e0) + 5 * c(e1) + 5 * c(e1)))
    circ = qcirc.Circuit(7, measurements=[(i, i) for i in range(7)], n_clbits=7)
    for _ in range(5):
        circ.cx(3, 0)
        circ.cx(6, 4)
    tree = [(0, 1), (1, 2), (2, 3), (1, 4), (2, 5), (5, 6)]
    out.append(dict(name="QubitRideAlong", n=7, cm=tree + [(b, a) for a, b in tree], q1=[e5] * 7,
                    cx=sym({(0, 1): e4, (1, 2): e3, (2, 3): e1, (1, 4): e1, (2, 5): e3, (5, 6): e4}),
                    circuit=circ, swaps=4,
                    fidelity=3 * c(e4) + 3 * c(e3) + 3 * c(e4) + 3 * c(e3) + 5 * c(e1) + 5 * c(e1)))
    circ = qcirc.Circuit(3, measurements=[(i, i) for i in range(3)], n_clbits=3)
    circ.gate("x", 0)
    circ.gate("x", 2)
    out.append(dict(name="SingleQubitsCompete", n=3, cm=path(3), q1=[0.8, 0.1, 0.8],
                    cx=sym({(0, 1): 0.1, (1, 2): 0.1}), circuit=circ, swaps=1,
                    fidelity=3 * c(0.1) + c(0.8) + c(0.1)))
    return out


FIDELITY_KNOWN_ANSWER_CONFIG = dict(layering="disjoint_2q_blocks", initial_layout="identity",
                                    heuristic="fidelity_best_location", lookahead_heuristic="none",
                                    pre_mapping_optimizations=False, post_mapping_optimizations=False,
                                    swap_on_first_layer=True)


def check_fidelity_known_answer(k, backend_for):
    """backend_for(ra) -> (backend, lib); routes k's circuit and checks the
    reference's expectations and the compiled reference's result."""
    ra = ref().arch_create(k["n"], sorted(set(k["cm"])))
    ra.set_error_rates(k["q1"], [k["cx"][tuple(int(x) for x in e)] for e in ra.edges])
    be, lib = backend_for(ra, k)
    msg, out, rr = run_pair(be, lib, ra, k["circuit"], **FIDELITY_KNOWN_ANSWER_CONFIG)
    assert msg == "", (k["name"], msg)
    assert rr.n_layers == 1 and rr.swaps == k["swaps"], k["name"]
    assert abs(rr.total_log_fidelity - k["fidelity"]) <= 1e-6, k["name"]
    assert len(out.swaps) == k["swaps"], k["name"]
    # one layer: the result node's fixed cost is the whole circuit's cost
    assert abs(out.layers[0]["cost_fixed"] - k["fidelity"]) <= 1e-6, k["name"]


# ---------------------------------------------------------------------------
# DataLogger parity (src/sc/DataLogger.cpp:216-269) and the reference's
# HeuristicProperties checks (test/sc/heuristic/test_heuristic.cpp:584-775)
# ---------------------------------------------------------------------------
# the fixture's settings (test_heuristic.cpp:261-268)
DATALOG_CONFIG = dict(debug=True, automatic_layer_splits=False, initial_layout="identity",
                      layering="disjoint_2q_blocks", lookahead_heuristic="none")
DATALOG_CIRCUITS = ["3_17_13", "ex-1_166", "ham3_102", "miller_11", "4gt11_84", "4mod5-v0_20", "mod5d1_63"]


def check_node_log(backend, lib, arch_file, calib, circuit, heuristic, node_log_capacity=200000):
    """Routes `circuit` with a node log and compares it, file by file and line
    by line, with what the compiled reference's DataLogger writes; then the
    reference's property checks on the log.  Returns the number of nodes."""
    import tempfile
    from qmap_b200 import datalog
    ra = ref().arch_load(os.path.join(ARCHS, arch_file), os.path.join(ARCHS, calib) if calib else None)
    kw = dict(DATALOG_CONFIG, heuristic=heuristic)
    path = os.path.join(CIRCUITS, circuit + ".qasm")
    d = tempfile.mkdtemp(prefix="qmap_datalog_") + os.sep
    rr = ref().map_qasm(ra, path, refapi.make_config(**kw), d)
    cfg = routing.make_config(**kw)
    circ = qcirc.load_qasm(path)
    out = routing.route_batch(backend, ra.n, cfg, [item_for(lib, ra.n, circ, cfg)],
                              want_layer_qubits=True, node_log_capacity=node_log_capacity)[0]
    assert out.status == 0
    assert compare(out, rr) == ""
    lines = datalog.layer_lines(out, rr.n_layers)
    total = 0
    tight = heuristic != 4  # isTight (Heuristic.hpp:96-110)
    for l in range(rr.n_layers):
        want = open(os.path.join(d, f"nodes_layer_{l}.csv")).read().splitlines()
        assert lines[l] == want, (circuit, arch_file, heuristic, l,
                                  next((a, b) for a, b in zip(lines[l] + [""], want + [""]) if a != b))
        total += len(want)
        # ---- test_heuristic.cpp:652-775 on the device's log
        nodes = {int(r["id"]): r for r in out.node_log if int(r["layer"]) == l}
        final = nodes[int(out.layers[l]["result_node_id"])]
        assert final["valid"] == 1
        if tight:
            assert abs(final["cost_heur"]) <= 1e-6   # tightness: goal nodes have h = 0
        assert final["lookahead_penalty"] == 0
        if heuristic in (0, 2, 3):  # isPrincipallyAdmissible (Heuristic.hpp:50-72)
            k = final
            while True:   # along the solution path f <= cost of the solution
                assert k["cost_fixed"] + k["cost_heur"] <= final["cost_fixed"] + 1e-6
                if k["root"]:
                    break
                k = nodes[int(k["parent"])]
    assert total == sum(int(x) for x in out.layers["generated_nodes"])
    return total
