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


def ref() -> refapi.Ref:
    global _ref
    if _ref is None:
        _ref = refapi.Ref()
    return _ref


def have_ref() -> bool:
    return os.path.exists(refapi.REF_LIB)


def identity_locations(n, circ) -> np.ndarray:
    loc = np.full(n, -1, dtype=np.int16)
    for q in circ.used_qubits():
        loc[q] = q
    return loc


def item_for(lib, n, circ, cfg):
    c, t = circ.gate_arrays()
    lc = routing.LayeredCircuit(lib, n, c, t, cfg.layering)
    init = identity_locations(n, circ) if cfg.initial_layout == 0 else None
    return routing.RouteItem(lc, init, cfg.iterative_bidirectional_routing_passes)


def compare(out, rr, rel=0.0):
    """out: routing.RouteOutput, rr: refapi.RefMapResult.  Returns '' if every
    final-pass aStarMap call agrees, else a description of the first
    difference.  rel > 0 allows that relative error on the four cost fields."""
    nl = rr.n_layers
    if (rr.call_flags & 2).any():
        return "reference split a layer (not comparable call by call)"
    if out.status != 0:
        return f"status {out.status}"
    if out.n_layers_done != nl:
        return f"layers done {out.n_layers_done} != {nl}"
    calls = rr.calls[len(rr.calls) - nl:]
    qa = rr.qubits_after[len(rr.calls) - nl:]
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
        c, t = circ.gate_arrays()
        rr = ref().map_gates(ra, circ.n_qubits, c, t, cfg_r)
    out = routing.route_batch(backend, ra.n, cfg, [item_for(lib, ra.n, circ, cfg)],
                              want_layer_qubits=True)[0]
    return compare(out, rr, rel), out, rr


def cuda_backend_for(ra, arch=None):
    """CudaBackend on a qmap_b200 Architecture equal to the reference arch."""
    from qmap_b200 import _abi
    from qmap_b200.architecture import Architecture
    if arch is None:
        arch = Architecture(ra.n, [tuple(e) for e in ra.edges])
    return routing.CudaBackend(_abi.load(), arch.handle), arch


def smoke_check() -> int:
    """Used by __graft_entry__.smoke(): configs 1 and 2 on cuda:0 vs the oracle."""
    from qmap_b200 import _abi
    lib = _abi.load()
    total = 0
    for arch_file, qasm, kw in (
            ("ibm_qx4.arch", "4gt13_92.qasm", dict(lookahead_heuristic="none")),
            ("ibm_qx5.arch", "qft_16.qasm", dict())):
        ra = ref().arch_load(os.path.join(ARCHS, arch_file))
        be, arch = cuda_backend_for(ra)
        for inc in (False, True):
            if not np.array_equal(arch.distance_table(inc), ra.distance_table(inc)):
                raise AssertionError(f"{arch_file}: distance table differs from the oracle")
        msg, out, rr = run_pair(be, lib, ra, os.path.join(CIRCUITS, qasm), **kw)
        if msg:
            raise AssertionError(f"{qasm} on {arch_file}: {msg}")
        total += rr.n_layers
    return total
