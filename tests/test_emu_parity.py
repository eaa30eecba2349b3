"""CPU suite: the HOST DEBUG BUILD of the routing kernel's source
(tests/emu/emu.cpp compiles mqt-qmap_b200/csrc/route_core.hpp with a 1-lane
group) against the committed outputs of the reference.  This exercises the
search logic the sm_100a kernel runs - candidate order, open-list dedup,
staged top selection, lookahead tables, fidelity scoring, placement - where
no GPU exists; the CUDA kernel itself is checked by the -m gpu tests.
Cost tables are inputs here (taken from the fixtures): the table kernels exist
only in CUDA."""
import os

import numpy as np
import pytest

import emu_backend
import parity
from qmap_b200 import _abi, routing

T = parity.golden_tables()
_backends = {}


def emu_for(name, calib=None):
    key = (name, calib)
    if key not in _backends:
        n, e = parity.read_arch_file(name)
        fid = None
        if calib:
            fid = (T[f"{name}_c1"], T[f"{name}_c2"], T[f"{name}_cswap"], T[f"{name}_fid"])
        _backends[key] = emu_backend.EmuBackend(n, e, T[f"{name}_dist"], T[f"{name}_dist_rev"],
                                                fid, pool_nodes=1 << 17)
    return _backends[key]


RUNS = [g for g in parity.golden_runs() if not g.meta["split"] and g.generated_nodes <= 50000]


@pytest.mark.parametrize("part", range(8))
def test_kernel_source_matches_reference_outputs(part):
    assert len(RUNS) > 230
    bad = []
    for g in RUNS[part::8]:
        be = emu_for(g.meta["arch"], g.meta["calibration"])
        msg, _ = parity.run_golden(be, be.lib, be.n, g)
        if msg:
            bad.append((g.label, msg))
    assert not bad, bad[:5]


def test_split_request_is_reported():
    """A search that hits the automatic split limit stops with
    QMG_SPLIT_LIMIT_HIT and names the layer (the host splits and resumes)."""
    g = next(g for g in parity.golden_runs() if g.meta["split"] and g.generated_nodes < 5000)
    be = emu_for(g.meta["arch"], g.meta["calibration"])
    msg, out = parity.run_golden(be, be.lib, be.n, g, backend_splits=False)
    assert out.status == _abi.QMG_SPLIT_LIMIT_HIT
    assert out.n_layers_done < g.n_layers


SPLIT_RUNS = [g for g in parity.golden_runs() if g.meta["split"] and g.generated_nodes <= 50000]


def test_layer_splits_match_reference_outputs():
    """Cases in which the reference split layers (final pass and pseudo-routing
    passes): with room for the new layers the search splits them itself
    (Mapper::splitLayer) and restarts on the uncleared open list like the
    reference - every call agrees bit for bit, stale nodes included."""
    assert len(SPLIT_RUNS) >= 28
    bad, splits = [], 0
    for g in SPLIT_RUNS:
        be = emu_for(g.meta["arch"], g.meta["calibration"])
        msg, out = parity.run_golden(be, be.lib, be.n, g)
        if msg:
            bad.append((g.label, msg))
        else:
            splits += len(out.split_layers)
            assert out.n_layers_done == g.n_layers
    assert not bad, bad[:3]
    assert splits > 100


def test_lean_outputs_match_full_records():
    """layer_n_swaps (compact per-layer output) equals the n_swaps of the full
    qmg_layer_result records"""
    g = next(g for g in RUNS if g.meta["circuit"] == "qft_16.qasm" and g.meta["arch"] == "ibm_qx5")
    be = emu_for(g.meta["arch"])
    from qmap_b200 import circuit as qcirc
    kw = dict(g.meta["config"])
    cfg = routing.make_config(**kw)
    circ = qcirc.load_qasm(os.path.join(parity.CIRCUITS, g.meta["circuit"]))
    item = parity.item_for(be.lib, be.n, circ, cfg)
    full = routing.route_batch(be, be.n, cfg, [item])[0]
    lean = routing.route_batch(be, be.n, cfg, [item], lean=True)[0]
    assert lean.layers is None
    assert np.array_equal(lean.layer_n_swaps, full.layers["n_swaps"])
    assert np.array_equal(lean.swaps, full.swaps)
    for l in (0, 5, len(full.layers) - 1):
        assert np.array_equal(lean.layer_swaps(l), full.layer_swaps(l))


@pytest.mark.parametrize("arch,calib", [("ibm_qx5", None), ("ibmq_tokyo", None),
                                        ("ibmq_london", "ibmq_london.csv")])
def test_random_circuits_with_layer_splits_against_oracle(arch, calib):
    """Seeded random circuits with tiny split limits: all five heuristics,
    pseudo-routing passes in both directions (splits there are undone after
    the pass), identity and dynamic layouts - records, SWAPs, layouts, the
    split log and the node totals equal the CPU restatement's."""
    from oracle import oracleapi
    from qmap_b200 import circuit as qcirc
    be = emu_for(arch, calib)
    n, e = parity.read_arch_file(arch)
    oa = oracleapi.OracleArch(n, e)
    if calib:
        oa.set_fidelity_costs(T[f"{arch}_c1"], T[f"{arch}_c2"], T[f"{arch}_cswap"])
    ob = oracleapi.OracleBackend(oa)
    splits = 0
    for seed in range(20):
        circ = qcirc.random_circuit(n - (seed % 3) if n > 6 else n, 20 + 2 * seed, seed=seed)
        heur = seed % 5 if calib else seed % 4
        cfg = routing.make_config(
            heuristic=heur, layering=("disjoint_qubits", "disjoint_2q_blocks")[seed % 2],
            initial_layout=("dynamic", "identity")[(seed // 2) % 2],
            lookahead_heuristic="none" if heur == 4 else
            ("gate_count_max_distance", "none", "gate_count_sum_distance")[seed % 3],
            iterative_bidirectional_routing_passes=(seed // 3) % 2,
            automatic_layer_splits_node_limit=(2, 3, 5, 9)[seed % 4], debug=True)
        item = parity.item_for(be.lib, n, circ, cfg)
        a = routing.route_batch(be, n, cfg, [item], want_layer_qubits=True, extra_layers=200)[0]
        b = routing.route_batch(ob, n, cfg, [item], want_layer_qubits=True, extra_layers=200)[0]
        assert parity.compare_outputs(a, b) == "", (arch, seed)
        assert np.array_equal(a.split_layers, b.split_layers), (arch, seed)
        assert (a.expanded_nodes, a.generated_nodes) == (b.expanded_nodes, b.generated_nodes)
        splits += len(b.split_layers)
    assert splits > 50


@pytest.mark.parametrize("env", [{"QMG_EMU_BLK": "4,8"}, {"QMG_EMU_BLK": "7,1"},
                                 {"QMG_EMU_BLK": "4,8,2", "QMG_EMU_START_FULL": "1"},
                                 {"QMG_EMU_BLK": "512,4096,1", "QMG_EMU_START_FULL": "1"},
                                 # (a call resumed after a layer split cannot move to the
                                 # large pool: the tiny private pool is for runs without splits)
                                 {"QMG_EMU_PRIVATE_POOL": "6", "QMG_EMU_BLK": "4,8,2"},
                                 {"QMG_EMU_PRIVATE_POOL": "40", "QMG_EMU_START_FULL": "1"},
                                 {"QMG_EMU_START_FULL": "1"},
                                 # other settings of the tuning macros: scan before the hash
                                 # table, 4 comparisons / open entries in flight, unrolled scans
                                 {"QMG_EMU_BLK": "4,8,2", "QMG_EMU_START_FULL": "1",
                                  "QMG_EMU_DEFS": "QMG_TABLE_FROM_START=0,QMG_FLIGHT=4,QMG_WIDE=4"},
                                 {"QMG_EMU_START_FULL": "1",
                                  "QMG_EMU_DEFS": "QMG_FLIGHT=1,QMG_WIDE=1,QMG_SCAN1_UNROLL=4"}])
def test_full_search_variant_matches_reference_outputs(env):
    """The kernel routes an instance with the lean search variant and again
    with the full one (per-block tops of the open list, layer splits) once an
    open list outgrows kBlkMin.  Built with blocks of 4 (7) entries from 8 (1)
    open nodes on, the golden cases run through the hand-over and the blocked
    open list; QMG_EMU_START_FULL=1 uses the full variant from the start."""
    import subprocess
    import sys
    code = (
        "import sys; sys.path[:0] = [%r, %r, %r]\n"
        "import emu_backend, parity, test_emu_parity as t\n"
        "import os\n"
        "runs = t.RUNS[::6] + t.SPLIT_RUNS[::3]\n"
        "if os.environ.get('QMG_TEST_NO_SPLITS'):\n"
        "    runs = [g for g in t.RUNS[::4] if 'splits_node_limit' not in g.label]\n"
        "bad = []\n"
        "for g in runs:\n"
        "    be = t.emu_for(g.meta['arch'], g.meta['calibration'])\n"
        "    msg, _ = parity.run_golden(be, be.lib, be.n, g)\n"
        "    if msg: bad.append((g.label, msg))\n"
        "assert len(runs) > 30 and not bad, bad[:3]\n"
        "print('ok', len(runs))\n"
    ) % (os.path.dirname(__file__), os.path.join(os.path.dirname(__file__), "emu"),
         os.path.join(parity.ROOT, "mqt-qmap_b200"))
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, **env),
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.startswith("ok")


@pytest.mark.parametrize("k", parity.fidelity_known_answers(), ids=lambda k: k["name"])
def test_reference_fidelity_known_answers(k):
    """test_heuristic.cpp:1769-1981 on the host build of the kernel source."""
    parity.check_fidelity_known_answer(
        k, lambda ra, _k: (lambda be: (be, be.lib))(emu_backend.EmuBackend.from_ref_arch(ra)))


@pytest.mark.parametrize("arch,calib,heuristics", [
    ("ibmq_yorktown.arch", None, [0, 1, 2, 3]),
    ("ibmq_london.arch", "ibmq_london.csv", [0, 3, 4]),
    ("ibm_qx5.arch", None, [0, 2]),
])
def test_node_log_equals_the_reference_datalogger(arch, calib, heuristics):
    """Row f3: every generated node of the HeuristicProperties fixture runs
    (test_heuristic.cpp:224-775), in the DataLogger's CSV schema, equal to the
    files the compiled reference writes."""
    ra = parity.ref().arch_load(os.path.join(parity.ARCHS, arch),
                                os.path.join(parity.ARCHS, calib) if calib else None)
    be = emu_backend.EmuBackend.from_ref_arch(ra, pool_nodes=1 << 17)
    total = 0
    for circuit in parity.DATALOG_CIRCUITS:
        for h in heuristics:
            total += parity.check_node_log(be, be.lib, arch, calib, circuit, h)
    assert total > 300
