"""-m gpu: the reference's own HeuristicMapper::map() run through the compiled
C++ drop-in (integration/GpuHeuristicMapper.hpp : HeuristicMapper, linked
against libqmap_b200.so; built by `make -C oracle dropin` into oracle/_ref/
libqmapdropin.so with the reference's sources) against plain HeuristicMapper
(oracle/_ref/libqmapref.so): per-call records, layouts, SWAPs, the emitted
circuit and the result counters.  Both libraries export the same driver API
(oracle/ref_driver.cpp), so refapi drives either."""
import os

import numpy as np
import pytest

import parity
from oracle import refapi

pytestmark = pytest.mark.gpu
DROPIN = os.path.join(parity.ROOT, "oracle", "_ref", "libqmapdropin.so")


@pytest.fixture(scope="module")
def dropin():
    if not os.path.exists(DROPIN):
        pytest.fail(f"{DROPIN} missing: run `make -C oracle dropin` where /root/reference exists")
    return refapi.Ref(DROPIN)


def _same(a, b, label):
    """a: through GpuHeuristicMapper, b: plain HeuristicMapper"""
    nl = b.n_layers
    assert a.n_layers == nl, label
    assert (a.swaps, a.direction_reverse, a.output_gates, a.output_cnots) == \
        (b.swaps, b.direction_reverse, b.output_gates, b.output_cnots), label
    # the final routing pass, call by call (the reference also records its
    # pseudo-routing calls; the drop-in runs those on the device)
    ka = np.nonzero((a.call_flags & 2) == 0)[0][-nl:] if nl else []
    kb = np.nonzero((b.call_flags & 2) == 0)[0][-nl:] if nl else []
    for k in ("cost_fixed", "cost_fixed_reversals", "cost_heur", "lookahead_penalty", "solution_depth",
              "result_node_id", "n_swaps", "expanded_nodes", "generated_nodes", "solution_nodes",
              "expanded_nodes_after_first_solution", "early_termination"):
        assert np.array_equal(a.calls[ka][k], b.calls[kb][k]), (label, k)
    assert np.array_equal(a.qubits_after[ka], b.qubits_after[kb]), label
    assert np.array_equal(a.locations_after[ka], b.locations_after[kb]), label
    for x, y in zip(ka, kb):
        ca, cb = a.calls[x], b.calls[y]
        assert np.array_equal(a.swap_pairs[ca["swap_offset"]: ca["swap_offset"] + ca["n_swaps"]],
                              b.swap_pairs[cb["swap_offset"]: cb["swap_offset"] + cb["n_swaps"]]), label
    assert np.array_equal(a.mapped_ops, b.mapped_ops), label          # the emitted circuit
    assert np.array_equal(a.initial_layout, b.initial_layout), label
    assert np.array_equal(a.output_permutation, b.output_permutation), label
    assert (a.expanded_nodes, a.generated_nodes) == (b.expanded_nodes, b.generated_nodes), label


def test_golden_runs_through_the_compiled_drop_in(dropin):
    """The golden configurations (all heuristics, layerings, layouts,
    lookaheads, bidirectional passes, early termination, fidelity-aware
    routing on ibmq_london, layer splits) through the reference's map()."""
    R = parity.ref()
    archs = {}
    n_runs = 0
    for g in parity.golden_runs():
        m = g.meta
        key = (m["arch"], m["calibration"])
        if key not in archs:
            files = (os.path.join(parity.ARCHS, m["arch"] + ".arch"),
                     os.path.join(parity.ARCHS, m["calibration"]) if m["calibration"] else None)
            archs[key] = (R.arch_load(*files), dropin.arch_load(*files))
        ra, da = archs[key]
        cfg = refapi.make_config(debug=True, **m["config"])
        path = os.path.join(parity.CIRCUITS, m["circuit"])
        _same(dropin.map_qasm(da, path, cfg), R.map_qasm(ra, path, cfg), g.label)
        n_runs += 1
    assert n_runs > 300


def test_drop_in_throws_like_the_reference(dropin):
    """InvalidCircuits (test_heuristic.cpp:951-960): cx(0,1) on a 2-qubit
    architecture without any edge makes the search run dry -> QMAPException
    "No viable mapping found." from either mapper."""
    c, t = np.array([0], dtype=np.int32), np.array([1], dtype=np.int32)
    for lib in (parity.ref(), dropin):
        a = lib.arch_create(2, [])
        with pytest.raises(RuntimeError, match="No viable mapping found"):
            lib.map_gates(a, 2, c, t, refapi.make_config())
