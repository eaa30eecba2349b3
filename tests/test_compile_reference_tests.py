"""The reference's own Python tests of the heuristic mapper
(test/python/test_heuristic_mapper.py:27-83), re-expressed for
qmap_b200.compile: the same circuits on the same 5-qubit backend, `timeout`
false, the SWAP count where the reference asserts one, and the mapped circuit
checked by the replay verifier (the reference uses mqt.qcec, which is not part
of it).  -m "not gpu": routed by the host build of the kernel source, cost
tables from the compiled reference."""
import numpy as np
import pytest

import emu_backend
import parity
import qmap_b200
from qmap_b200 import architecture as ar
from qmap_b200 import circuit as qcirc
from qmap_b200 import mapper

COUPLING = [(0, 1), (1, 0), (1, 2), (2, 1), (1, 3), (3, 1), (3, 4), (4, 3)]  # GenericBackendV2 of the fixture


def emu_for(a):
    """The host build of the kernel source for Architecture `a`: cost tables
    from the compiled reference built on the same coupling map / error rates."""
    ra = parity.ref().arch_create(a.n_qubits, a.coupling_map)
    fid = None
    if a.fidelity_available:
        p = a.properties
        q1 = [p.average_single_qubit_error_rate(q) if q in p.single_qubit_error else np.nan
              for q in range(a.n_qubits)]
        cx = [p.two_qubit_error.get(e, np.nan) for e in a.coupling_map]
        ra.set_error_rates(q1, cx)
        fid = (ra.single_qubit_fidelity_costs(), ra.two_qubit_fidelity_costs(), ra.swap_fidelity_costs(),
               ra.fidelity_distance_tables())
        assert np.array_equal(fid[0], a.single_qubit_fidelity_costs())   # (host -log2 costs, bit for bit)
        assert np.array_equal(fid[1], a.two_qubit_fidelity_costs())
        assert np.array_equal(fid[2], a.swap_fidelity_costs())
    return emu_backend.EmuBackend(a.n_qubits, np.array(a.coupling_map, dtype=np.uint16).reshape(-1, 2),
                                  ra.distance_table(False), ra.distance_table(True), fid, pool_nodes=1 << 14)


@pytest.fixture(scope="module")
def backend():
    a = ar.Architecture(5, COUPLING)
    return a, emu_for(a)


def _circuit(n, gates):
    c = qcirc.Circuit(n, n_clbits=n)
    c.h(0)
    for a, b in gates:
        c.cx(a, b)
    c.measurements = [(q, q) for q in range(n)]  # measure_all
    return c


def _compile_and_verify(backend, circ):
    a, be = backend
    mapped, results = qmap_b200.compile(circ, arch=a, lib=be.lib, backend=be)
    assert results.timeout is False
    m = mapper.HeuristicMapper(circ, a, lib=be.lib, backend=be)
    m.map()
    assert mapper.verify_mapping(circ, m) == ""
    assert [(o.name, o.qubits) for o in m.mapped_ops] == [(o.name, o.qubits) for o in mapped.ops]
    return results


def test_heuristic_no_swaps_trivial_layout(backend):
    _compile_and_verify(backend, _circuit(3, [(0, 1), (1, 2)]))


def test_heuristic_no_swaps_non_trivial_layout(backend):
    _compile_and_verify(backend, _circuit(4, [(0, 1), (0, 2), (0, 3)]))


def test_heuristic_non_trivial_swaps(backend):
    results = _compile_and_verify(backend, _circuit(3, [(0, 1), (1, 2), (2, 0)]))
    assert results.output_swaps == 1


# ---------------------------------------------------------------------------
# test/python/test_compile.py:36-258 (the heuristic method)
# ---------------------------------------------------------------------------
def example_circuit():
    return _circuit(3, [(0, 1), (1, 2)])


def _compile_on(arch, calibration=None, **kw):
    a = qmap_b200.compile.__globals__["load_architecture"](arch)
    qmap_b200.compile.__globals__["load_calibration"](a, calibration)
    be = emu_for(a)
    circ = example_circuit()
    mapped, results = qmap_b200.compile(circ, arch=a, lib=be.lib, backend=be, **kw)
    assert results.timeout is False
    m = mapper.HeuristicMapper(circ, a, lib=be.lib, backend=be)
    m.map(results.config)
    assert mapper.verify_mapping(circ, m) == ""
    return mapped, results


@pytest.mark.parametrize("arch", ["IBM_QX4", "IBM_QX5", "IBMQ_Yorktown", "IBMQ_London", "IBMQ_Bogota",
                                  "IBMQ_Tokyo", "Rigetti_Agave", "Rigetti_Aspen"])
def test_available_architectures_str(arch):
    _compile_on(arch)


def test_architecture_from_file(tmp_path):
    f = tmp_path / "test_architecture.arch"
    f.write_text("3\n0 1\n0 2\n1 2\n")
    _compile_on(str(f))


def test_architecture_from_python():
    _compile_on(ar.Architecture(3, {(0, 1), (0, 2), (1, 2)}))


def test_calibration_from_file(tmp_path):
    f = tmp_path / "test_calibration.cal"
    f.write_text('Header\nQ0,0,0,0,1e-2,1e-4,"0_1: 1e-2, 0_2: 1e-2"\nQ1,0,0,0,1e-2,1e-4,"1_2: 1e-2"\n'
                 "Q2,0,0,0,1e-2,1e-4, \n")
    _, results = _compile_on(None, calibration=str(f))
    assert results.output_qubits == 3


def test_parameters():
    from qmap_b200 import _abi
    properties = ar.Properties()
    for q in range(3):
        properties.set_single_qubit_error_rate(q, "x", 0.01)
    for a, b in ((0, 1), (1, 0), (1, 2), (2, 1)):
        properties.set_two_qubit_error_rate(a, b, 0.02)
    arch = ar.Architecture(3, {(0, 1), (1, 0), (1, 2), (2, 1)}, properties)
    _, results = _compile_on(
        arch, method="heuristic", heuristic="gate_count_max_distance", initial_layout="dynamic",
        iterative_bidirectional_routing_passes=1, layering="individual_gates",
        automatic_layer_splits_node_limit=5000, lookahead_heuristic="gate_count_max_distance",
        lookaheads=15, lookahead_factor=0.5, pre_mapping_optimizations=True,
        post_mapping_optimizations=True, verbose=True, debug=True)
    c = results.config
    assert c.heuristic == _abi.HEURISTICS["gate_count_max_distance"]
    assert c.lookahead_heuristic == _abi.LOOKAHEADS["gate_count_max_distance"]
    assert c.initial_layout == _abi.LAYOUTS["dynamic"]
    assert c.iterative_bidirectional_routing_passes == 1
    assert c.layering == _abi.LAYERINGS["individual_gates"]
    assert c.automatic_layer_splits == 1 and c.automatic_layer_splits_node_limit == 5000
    assert c.nr_lookaheads == 15 and c.lookahead_factor == 0.5
    assert c.pre_mapping_optimizations == 1 and c.post_mapping_optimizations == 1
    assert c.verbose == 1 and c.debug == 1
    _, results = _compile_on(
        arch, method="heuristic", heuristic="fidelity_best_location", initial_layout="identity",
        iterative_bidirectional_routing_passes=None, layering="disjoint_qubits",
        automatic_layer_splits_node_limit=None, lookahead_heuristic=None,
        pre_mapping_optimizations=False, post_mapping_optimizations=False, verbose=False, debug=False)
    c = results.config
    assert c.heuristic == _abi.HEURISTICS["fidelity_best_location"]
    assert c.lookahead_heuristic == _abi.LOOKAHEADS["none"]
    assert c.initial_layout == _abi.LAYOUTS["identity"]
    assert c.iterative_bidirectional_routing_passes == 0
    assert c.layering == _abi.LAYERINGS["disjoint_qubits"]
    assert c.automatic_layer_splits == 0 and c.nr_lookaheads == 0
    assert c.pre_mapping_optimizations == 0 and c.post_mapping_optimizations == 0
    assert c.verbose == 0 and c.debug == 0


def test_layer_splitting_known_answer():
    """HeuristicTestFidelity.LayerSplitting (test/sc/heuristic/test_heuristic.cpp:
    1983-2141): a 3 x 4 grid with hand-picked error rates, one layer of 50 + 2
    X gates and two CX blocks; a node limit of 1 forces splits until the layers
    cannot be split any more.  Expected by the reference: 5 layers, 8 SWAPs and
    the total log-fidelity written out below."""
    import math
    cm = set()
    for r in range(3):
        for c in range(4):
            q = 4 * r + c
            if c < 3:
                cm |= {(q, q + 1), (q + 1, q)}
            if r < 2:
                cm |= {(q, q + 4), (q + 4, q)}
    e5, e4, e3, e2, e1, e0 = 0.99, 0.9, 0.5, 0.4, 0.1, 0.01
    props = ar.Properties()
    for q, e in enumerate([e5, e5, e5, e5, e3, e3, e5, e3, e2, e1, e5, e2]):
        props.set_single_qubit_error_rate(q, "x", e)
    two = {(0, 1): e4, (1, 2): e4, (2, 3): e1, (0, 4): e5, (1, 5): e5, (2, 6): e4, (3, 7): e5,
           (4, 5): e3, (5, 6): e5, (6, 7): e5, (4, 8): e0, (5, 9): e3, (6, 10): e1, (7, 11): e3,
           (8, 9): e5, (9, 10): e4, (10, 11): e5}
    for (a, b), e in two.items():
        props.set_two_qubit_error_rate(a, b, e)
        props.set_two_qubit_error_rate(b, a, e)
    a = ar.Architecture(12, cm, props)
    be = emu_for(a)
    circ = qcirc.Circuit(12, n_clbits=12)
    for _ in range(50):
        circ.gate("x", 4)
    circ.gate("x", 5)
    circ.gate("x", 7)
    for _ in range(5):
        circ.cx(3, 0)
        circ.cx(9, 2)
    circ.measurements = [(q, q) for q in range(12)]
    m = mapper.HeuristicMapper(circ, a, lib=be.lib, backend=be)
    res = m.map(layering="disjoint_2q_blocks", initial_layout="identity", heuristic="fidelity_best_location",
                lookahead_heuristic="none", pre_mapping_optimizations=False, post_mapping_optimizations=False,
                swap_on_first_layer=True, automatic_layer_splits=True, automatic_layer_splits_node_limit=1)
    assert mapper.verify_mapping(circ, m) == ""
    assert res.input_layers == 5          # originally 1, split into 5 during A*
    assert res.output_swaps == 8
    c4, c3, c2, c1, c0 = (-math.log2(1 - e) for e in (e4, e3, e2, e1, e0))
    expected = 3 * (c3 + c3 + c3 + c4 + c4 + c0 + c4 + c4) + 50 * c1 + c3 + c2 + 5 * c1 + 5 * c1
    assert abs(res.total_log_fidelity - expected) < 1e-6


def test_early_termination_known_answers():
    """Functionality.earlyTermination (test_heuristic.cpp:1283-1362): x(0),
    cx(1,2) on a 7-qubit line with calibration, fidelity-aware, limit 4: the
    first layer's benchmark record under every termination strategy."""
    cm = set()
    for q in range(6):
        cm |= {(q, q + 1), (q + 1, q)}
    props = ar.Properties()
    for q, e in enumerate([0.9, 0.5, 0.5, 0.5, 0.5, 0.5, 0.1]):
        props.set_single_qubit_error_rate(q, "x", e)
    for a, b in cm:
        props.set_two_qubit_error_rate(a, b, 0.01)
    arch = ar.Architecture(7, cm, props)
    be = emu_for(arch)
    circ = qcirc.Circuit(7, n_clbits=7)
    circ.gate("x", 0)
    circ.cx(1, 2)
    circ.measurements = [(q, q) for q in range(7)]
    base = dict(layering="disjoint_qubits", initial_layout="identity", automatic_layer_splits=False,
                debug=True, heuristic="fidelity_best_location", lookahead_heuristic="none",
                early_termination_limit=4)

    def first_layer(strategy):
        m = mapper.HeuristicMapper(circ, arch, lib=be.lib, backend=be)
        res = m.map(early_termination=strategy, **base)
        assert mapper.verify_mapping(circ, m) == ""
        return res.layer_benchmark[0]

    assert not first_layer("none")["early_termination"]
    for strategy, field in (("expanded_nodes_after_first_solution", "expanded_nodes_after_first_solution"),
                            ("expanded_nodes_after_current_optimal_solution", "expanded_nodes_after_optimal_solution"),
                            ("expanded_nodes", "expanded_nodes"),
                            ("solution_nodes", "solution_nodes"),
                            ("solution_nodes_after_current_optimal_solution",
                             "solution_nodes_after_optimal_solution")):
        r = first_layer(strategy)
        assert r["early_termination"] and r[field] == 4, strategy


def test_invalid_settings():
    """Functionality.InvalidSettings (test_heuristic.cpp:895-916)."""
    props = ar.Properties()
    props.set_single_qubit_error_rate(0, "x", 0.1)
    arch = ar.Architecture(1, [], props)
    be = emu_for(ar.Architecture(1, []))
    circ = qcirc.Circuit(1)
    circ.gate("x", 0)
    m = mapper.HeuristicMapper(circ, arch, lib=be.lib, backend=be)
    for bad in (dict(layering="odd_gates"), dict(layering="qubit_triangle"),
                dict(layering="individual_gates", heuristic="fidelity_best_location")):
        with pytest.raises(qmap_b200.QmapError):
            m.map(heuristic=bad.pop("heuristic", "gate_count_max_distance"),
                  lookahead_heuristic="gate_count_max_distance", **bad)


@pytest.mark.parametrize("layering,layers", [("disjoint_2q_blocks", 2), ("disjoint_qubits", 3),
                                             ("individual_gates", 6)])
def test_layering_known_answers(layering, layers):
    """LayeringTest (test_heuristic.cpp:1450-1543): x(0) x(1) cx(0,1) cx(2,3)
    cx(1,2) x(3) on a directed 4-qubit line; the number of layers per
    strategy, and one barrier between consecutive layers (the reference's
    count is one higher: mqt-core puts another barrier before the final
    measurements)."""
    arch = ar.Architecture(4, [(0, 1), (1, 2), (2, 3)])
    be = emu_for(arch)
    circ = qcirc.Circuit(4, n_clbits=4)
    circ.gate("x", 0)
    circ.gate("x", 1)
    circ.cx(0, 1)
    circ.cx(2, 3)
    circ.cx(1, 2)
    circ.gate("x", 3)
    circ.measurements = [(q, q) for q in range(3)]
    m = mapper.HeuristicMapper(circ, arch, lib=be.lib, backend=be)
    res = m.map(layering=layering, initial_layout="dynamic", pre_mapping_optimizations=False,
                post_mapping_optimizations=False, add_measurements_to_mapped_circuit=True,
                add_barriers_between_layers=True, automatic_layer_splits=False)
    assert res.input_layers == layers
    assert sum(1 for o in m.mapped_ops if o.name == "barrier") == layers - 1
    assert mapper.verify_mapping(circ, m) == ""


def test_heuristic_benchmark_known_answers():
    """Functionality.HeuristicBenchmark (test_heuristic.cpp:777-856): three CX
    on a 5-ring, identity layout, no lookahead - node counts, depths and
    branching factors of both layers and of the run."""
    cm = [(0, 1), (1, 0), (1, 2), (2, 1), (2, 3), (3, 2), (3, 4), (4, 3), (4, 0), (0, 4)]
    arch = ar.Architecture(5, cm)
    be = emu_for(arch)
    circ = qcirc.Circuit(5, n_clbits=5)
    circ.cx(4, 2)
    circ.cx(3, 1)
    circ.cx(4, 1)
    circ.measurements = [(q, q) for q in range(5)]
    m = mapper.HeuristicMapper(circ, arch, lib=be.lib, backend=be)
    res = m.map(heuristic="gate_count_max_distance", layering="disjoint_qubits", initial_layout="identity",
                pre_mapping_optimizations=False, post_mapping_optimizations=False,
                lookahead_heuristic="none", debug=True)
    lb = res.layer_benchmark
    assert len(lb) == 2
    assert (lb[0]["solution_depth"], lb[0]["generated_nodes"], lb[0]["expanded_nodes"]) == (1, 6, 1)
    assert (lb[1]["solution_depth"], lb[1]["generated_nodes"], lb[1]["expanded_nodes"]) == (1, 5, 1)
    for r, avg in ((lb[0], 5.0), (lb[1], 4.0)):
        assert (int(r["generated_nodes"]) - 1) / int(r["expanded_nodes"]) == avg
        assert abs(mapper.effective_branching_rate(int(r["expanded_nodes"]) + 1, int(r["solution_depth"])) - 1.0) < 1e-10
    assert (res.generated_nodes, res.expanded_nodes) == (11, 2)
    assert abs(res.average_branching_factor - 4.5) < 1e-10
    assert abs(res.effective_branching_factor - 1.0) < 1e-10


def test_benchmark_generated_nodes():
    """Functionality.BenchmarkGeneratedNodes (test_heuristic.cpp:869-893):
    cx(0,6) on IBM QX5 generates 30 nodes."""
    a = ar.Architecture().load_coupling_map("IBM_QX5")
    be = emu_for(a)
    circ = qcirc.Circuit(16, n_clbits=16)
    circ.cx(0, 6)
    circ.measurements = [(q, q) for q in range(16)]
    m = mapper.HeuristicMapper(circ, a, lib=be.lib, backend=be)
    res = m.map(heuristic="gate_count_max_distance", lookahead_heuristic="none", layering="individual_gates",
                automatic_layer_splits=False, initial_layout="identity", pre_mapping_optimizations=False,
                post_mapping_optimizations=False, debug=True)
    assert res.generated_nodes == 30 and res.layer_benchmark[0]["generated_nodes"] == 30
