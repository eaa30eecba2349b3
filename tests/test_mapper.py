"""CPU suite: the host mirror of HeuristicMapper (qmap_b200/mapper.py) - initial
layouts, split-and-resume, circuit emission, verifier - driven through the
host debug build of the kernel source; SWAP counts are checked against the
committed outputs of the reference."""
import os

import numpy as np
import pytest

import emu_backend
import parity
from qmap_b200 import _abi, mapper, routing
from qmap_b200 import architecture as ar
from qmap_b200 import circuit as qcirc

T = parity.golden_tables()
_cache = {}


def arch_and_backend(name):
    if name not in _cache:
        n, e = parity.read_arch_file(name)
        a = ar.Architecture(n, [tuple(int(v) for v in x) for x in e])
        be = emu_backend.EmuBackend(n, e, T[f"{name}_dist"], T[f"{name}_dist_rev"],
                                    pool_nodes=1 << 17)
        _cache[name] = (a, be)
    return _cache[name]


CASES = [g for g in parity.golden_runs()
         if g.meta["calibration"] is None and g.generated_nodes <= 20000][::3]


def test_mapped_circuits_verify_and_match_reference_swap_counts():
    assert len(CASES) > 60
    bad = []
    n_split = 0
    n_rerouted = 0
    for g in CASES:
        a, be = arch_and_backend(g.meta["arch"])
        circ = qcirc.load_qasm(os.path.join(parity.CIRCUITS, g.meta["circuit"]))
        m = mapper.HeuristicMapper(circ, a, lib=be.lib, backend=be)
        res = m.map(**g.meta["config"])
        msg = mapper.verify_mapping(circ, m)
        if msg:
            bad.append((g.label, msg))
            continue
        if g.meta["split"]:
            # layers are split by the library exactly like the reference
            # (pseudo-routing passes included): same final layer structure
            n_split += 1
            if res.rerouted_with_host_splits:
                # the reference's own result for this case places a CNOT on
                # qubits that are not coupled (see HeuristicMapper.map)
                n_rerouted += 1
                continue
            if res.input_layers != g.n_layers:
                bad.append((g.label, f"layers {res.input_layers} != {g.n_layers}"))
        if res.output_swaps != g.swaps:
            # (swap_on_first_layer is off: the first layer's SWAPs fold into
            # the initial layout in both implementations)
            bad.append((g.label, f"swaps {res.output_swaps} != {g.swaps}"))
    assert not bad, bad[:5]
    assert n_split >= 5
    assert n_rerouted <= 1


def test_static_and_identity_layouts():
    a, be = arch_and_backend("ibm_qx5")
    circ = qcirc.load_qasm(os.path.join(parity.CIRCUITS, "rd32-v0_66.qasm"))
    for layout in ("identity", "static", "dynamic"):
        m = mapper.HeuristicMapper(circ, a, lib=be.lib, backend=be)
        res = m.map(initial_layout=layout)
        assert mapper.verify_mapping(circ, m) == ""
        # Mapper::countGates: IBM QX5 is unidirectional - a SWAP is 3 CNOTs + 4 H,
        # a reversed CNOT 4 H
        assert res.output_cnots == res.input_cnots + 3 * res.output_swaps
        assert res.output_gates == (res.input_gates + 7 * res.output_swaps
                                    + 4 * res.output_direction_reverse)
        out = m.mapped_circuit()
        again = qcirc.parse_qasm(out.to_qasm())
        assert len(again.ops) == len(out.ops)


def test_parameter_checks_use_the_reference_messages():
    a, be = arch_and_backend("ibm_qx4")
    circ = qcirc.Circuit(2)
    circ.cx(0, 1)
    m = mapper.HeuristicMapper(circ, a, lib=be.lib, backend=be)
    with pytest.raises(_abi.QmapError, match="not suitable for heuristic mapper"):
        m.map(layering="odd_gates")
    with pytest.raises(_abi.QmapError, match="calibration data"):
        m.map(heuristic="fidelity_best_location", lookahead_heuristic="none")


def test_verifier_rejects_a_wrong_circuit():
    a, be = arch_and_backend("ibm_qx5")
    circ = qcirc.load_qasm(os.path.join(parity.CIRCUITS, "4mod5-v0_20.qasm"))
    m = mapper.HeuristicMapper(circ, a, lib=be.lib, backend=be)
    m.map()
    assert mapper.verify_mapping(circ, m) == ""
    k = next(i for i, o in enumerate(m.mapped_ops) if o.name == "swap")
    del m.mapped_ops[k]
    assert mapper.verify_mapping(circ, m) != ""


def test_compile_mirrors_the_reference_entry_point():
    """qmap_b200.compile (python/mqt/qmap/compile.py:59-171): argument names
    and defaults of the reference, SWAP counts of its committed outputs."""
    import inspect
    import qmap_b200
    sig = inspect.signature(qmap_b200.compile)
    want = ["circ", "arch", "calibration", "method", "heuristic", "initial_layout",
            "iterative_bidirectional_routing_passes", "layering", "automatic_layer_splits_node_limit",
            "early_termination", "early_termination_limit", "lookahead_heuristic", "lookaheads",
            "lookahead_factor", "encoding", "commander_grouping", "swap_reduction", "swap_limit",
            "include_WCNF", "use_subsets", "subgraph", "pre_mapping_optimizations",
            "post_mapping_optimizations", "add_measurements_to_mapped_circuit",
            "add_barriers_between_layers", "verbose", "debug", "visualizer"]
    assert list(sig.parameters)[:len(want)] == want
    assert sig.parameters["automatic_layer_splits_node_limit"].default == 5000
    assert sig.parameters["lookaheads"].default == 15 and sig.parameters["lookahead_factor"].default == 0.5
    with pytest.raises(ValueError, match="Either arch or calibration must be specified"):
        qmap_b200.compile("qreg q[1];", None)
    a, be = arch_and_backend("ibm_qx5")
    with pytest.raises(_abi.QmapError, match="not part of this library"):
        qmap_b200.compile("qreg q[1];", a, method="exact", lib=be.lib, backend=be)
    n = 0
    for g in CASES:
        cfg = g.meta["config"]
        if g.meta["arch"] != "ibm_qx5" or g.meta["split"] or set(cfg) - {"heuristic", "lookahead_heuristic",
                                                                        "layering", "initial_layout"}:
            continue
        path = os.path.join(parity.CIRCUITS, g.meta["circuit"])
        kw = dict(cfg)
        if kw.get("lookahead_heuristic") == "none":
            kw["lookahead_heuristic"] = None
        mapped, res = qmap_b200.compile(path, a, lib=be.lib, backend=be, **kw)
        assert res.output_swaps == g.swaps, g.label
        assert mapped.n_qubits == 16 and len(mapped.ops) == res.output_gates - 6 * res.output_swaps
        n += 1
    assert n >= 5


def _backend_with_calibration(name, calib):
    n, e = parity.read_arch_file(name)
    a = ar.Architecture().load_coupling_map(os.path.join(parity.ARCHS, name + ".arch"))
    fid = None
    if calib:
        a.load_properties(os.path.join(parity.ARCHS, calib))
        fid = (T[f"{name}_c1"], T[f"{name}_c2"], T[f"{name}_cswap"], T[f"{name}_fid"])
    be = emu_backend.EmuBackend(n, e, T[f"{name}_dist"], T[f"{name}_dist_rev"], fid, pool_nodes=1 << 17)
    return a, be


def test_mapping_results_json_equals_the_reference():
    """MappingResults::json() of the mirror against the compiled reference's:
    circuit / mapped_circuit counts (Mapper::countGates), Configuration::json(),
    swaps, layers, total (log) fidelity, the debug benchmark with average and
    effective branching factors - everything but the two timing fields."""
    import json
    from oracle import refapi
    R = parity.ref()
    cases = [g for g in parity.golden_runs() if not g.meta["split"] and g.generated_nodes <= 20000]
    picked = [g for g in cases if g.meta["calibration"]][::6] + [g for g in cases if not g.meta["calibration"]][::9]
    assert len(picked) > 25
    cache = {}
    for g in picked:
        m = g.meta
        key = (m["arch"], m["calibration"])
        if key not in cache:
            files = (os.path.join(parity.ARCHS, m["arch"] + ".arch"),
                     os.path.join(parity.ARCHS, m["calibration"]) if m["calibration"] else None)
            cache[key] = _backend_with_calibration(*key) + (R.arch_load(*files),)
        a, be, ra = cache[key]
        path = os.path.join(parity.CIRCUITS, m["circuit"])
        want = json.loads(R.map_qasm(ra, path, refapi.make_config(debug=True, **m["config"])).json)
        circ = qcirc.load_qasm(path)
        mp = mapper.HeuristicMapper(circ, a, lib=be.lib, backend=be)
        got = mp.map(debug=True, **m["config"]).json()
        for j in (got, want):
            j["statistics"].pop("mapping_time")
            j["statistics"]["benchmark"].pop("seconds_per_node")
        assert got == want, (g.label, got, want)


@pytest.mark.parametrize("arch,calib,heuristic", [
    ("ibm_qx5", None, "gate_count_max_distance"),
    ("ibmq_london", "ibmq_london.csv", "fidelity_best_location")])
def test_data_logging_directory_equals_the_reference(arch, calib, heuristic, tmp_path):
    """compile(..., visualizer=...) leaves what the reference's DataLogger
    leaves (DataLogger.cpp): nodes_layer_<i>.csv line by line, layer_<i>.json,
    architecture.json and mapping_result.json value by value (timing fields
    aside); input.qasm / output.qasm exist (mqt-core's text is not pinned)."""
    import json
    import types
    import qmap_b200
    from oracle import refapi
    R = parity.ref()
    a, be = _backend_with_calibration(arch, calib)
    a._dist, a._dist_rev = T[f"{arch}_dist"].copy(), T[f"{arch}_dist_rev"].copy()  # (no GPU here)
    if calib:
        a._fid = T[f"{arch}_fid"].copy()
    files = (os.path.join(parity.ARCHS, arch + ".arch"), os.path.join(parity.ARCHS, calib) if calib else None)
    ra = R.arch_load(*files)
    for circuit in parity.DATALOG_CIRCUITS[:3]:
        path = os.path.join(parity.CIRCUITS, circuit + ".qasm")
        kw = dict(parity.DATALOG_CONFIG, heuristic=heuristic)
        d_ref = str(tmp_path / (circuit + "_ref")) + os.sep
        d_got = str(tmp_path / (circuit + "_got")) + os.sep
        R.map_qasm(ra, path, refapi.make_config(**kw), d_ref)
        ckw = {k: v for k, v in kw.items() if k != "automatic_layer_splits"}
        ckw["lookahead_heuristic"] = None
        ckw["automatic_layer_splits_node_limit"] = None
        qmap_b200.compile(path, a, visualizer=types.SimpleNamespace(data_logging_path=d_got),
                          lib=be.lib, backend=be, **ckw)
        names = sorted(os.listdir(d_ref))
        assert sorted(os.listdir(d_got)) == names
        for f in names:
            if f.endswith(".csv"):
                assert open(d_got + f).read() == open(d_ref + f).read(), (circuit, f)
            elif f.endswith(".json"):
                x, y = json.load(open(d_got + f)), json.load(open(d_ref + f))
                if f == "mapping_result.json":
                    for j in (x, y):
                        j["statistics"].pop("mapping_time")
                        j["statistics"]["benchmark"].pop("seconds_per_node")
                        for lb in j["statistics"]["benchmark"]["layers"]:
                            lb.pop("seconds_per_node")
                assert x == y, (circuit, f)
