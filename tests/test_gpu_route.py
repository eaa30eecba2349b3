"""-m gpu: the routing kernel against the compiled reference (oracle/_ref),
call by call (costs, counters, result node, layout, SWAPs), through the C ABI."""
import glob
import os

import numpy as np
import pytest

import parity
from qmap_b200 import _abi, routing
from qmap_b200 import architecture as ar
from qmap_b200 import circuit as qcirc

pytestmark = pytest.mark.gpu


def _pair(arch_file, calib=None):
    ra = parity.ref().arch_load(os.path.join(parity.ARCHS, arch_file),
                                os.path.join(parity.ARCHS, calib) if calib else None)
    a = ar.Architecture().load_coupling_map(os.path.join(parity.ARCHS, arch_file))
    if calib:
        a.load_properties(os.path.join(parity.ARCHS, calib))
    be = routing.CudaBackend(_abi.load(), a.handle)
    return ra, a, be


def test_config1_4gt13_92_on_qx4():
    ra, a, be = _pair("ibm_qx4.arch")
    msg, out, rr = parity.run_pair(be, _abi.load(), ra, os.path.join(parity.CIRCUITS, "4gt13_92.qasm"),
                                   lookahead_heuristic="none")
    assert msg == ""
    assert (out.expanded_nodes, out.generated_nodes, len(out.swaps)) == (6, 90, 6)


def test_config2_qft16_on_qx5():
    ra, a, be = _pair("ibm_qx5.arch")
    msg, out, rr = parity.run_pair(be, _abi.load(), ra, os.path.join(parity.CIRCUITS, "qft_16.qasm"))
    assert msg == ""
    assert (out.expanded_nodes, out.generated_nodes, len(out.swaps)) == (75, 937, 75)


SMALL = sorted(glob.glob(os.path.join(parity.CIRCUITS, "*.qasm")), key=os.path.getsize)


@pytest.mark.parametrize("arch_file", ["ibm_qx5.arch", "ibmq_tokyo.arch"])
@pytest.mark.parametrize("heuristic", [0, 1, 2, 3])
@pytest.mark.parametrize("layering", ["individual_gates", "disjoint_qubits", "disjoint_2q_blocks"])
def test_sweep_gate_count(arch_file, heuristic, layering):
    ra, a, be = _pair(arch_file)
    bad = []
    for il in ("dynamic", "identity"):
        for la in ("none", "gate_count_max_distance", "gate_count_sum_distance"):
            for f in SMALL[:14:2]:
                msg, _, _ = parity.run_pair(be, _abi.load(), ra, f, heuristic=heuristic,
                                            layering=layering, initial_layout=il,
                                            lookahead_heuristic=la)
                if msg and "split" not in msg:
                    bad.append((os.path.basename(f), il, la, msg))
    assert not bad, bad[:5]


@pytest.mark.parametrize("layering", ["individual_gates", "disjoint_qubits", "disjoint_2q_blocks"])
def test_fidelity_london(layering):
    ra, a, be = _pair("ibmq_london.arch", "ibmq_london.csv")
    bad = []
    for f in SMALL:
        c = qcirc.load_qasm(f)
        if c.n_qubits > 5:
            continue
        for il in ("dynamic", "identity"):
            msg, _, _ = parity.run_pair(be, _abi.load(), ra, f, rel=1e-9, heuristic="fidelity_best_location",
                                        lookahead_heuristic="none", layering=layering, initial_layout=il)
            if msg and "split" not in msg:
                bad.append((os.path.basename(f), il, msg))
    assert not bad, bad[:5]


def test_bidirectional_passes():
    ra, a, be = _pair("ibmq_tokyo.arch")
    for f in SMALL[:8]:
        msg, _, _ = parity.run_pair(be, _abi.load(), ra, f, iterative_bidirectional_routing_passes=2)
        assert msg == "", (f, msg)


def test_batch_of_relabelled_instances_matches_single_runs():
    """Instances of one batch are independent: a batch of 64 relabelled copies
    of qft_16 gives, per instance, what the reference gives for that copy."""
    ra, a, be = _pair("ibm_qx5.arch")
    lib = _abi.load()
    base = qcirc.load_qasm(os.path.join(parity.CIRCUITS, "qft_16.qasm"))
    cfg = routing.make_config(initial_layout="identity")
    rng = np.random.default_rng(0)
    circs = [base.relabel(rng.permutation(16)) for _ in range(64)]
    items = [parity.item_for(lib, 16, c, cfg) for c in circs]
    outs = routing.route_batch(be, 16, cfg, items, want_layer_qubits=True)
    cfg_r = parity.refapi.make_config(initial_layout="identity", debug=True)
    for c, out in zip(circs[::8], outs[::8]):
        cc, tt = c.gate_arrays()
        rr = parity.ref().map_gates(ra, 16, cc, tt, cfg_r)
        assert parity.compare(out, rr) == ""
