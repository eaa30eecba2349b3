"""-m gpu: the table kernels (K1 blocked Floyd-Warshall, K1' replay of the
reference's Dijkstra, K2 single-edge-skip, K2' k-edge-skip) against the compiled
reference (oracle/_ref), through the C ABI."""
import ctypes as C
import glob
import os

import numpy as np
import pytest

import parity
from qmap_b200 import _abi
from qmap_b200 import architecture as ar

pytestmark = pytest.mark.gpu
DBL_MAX = np.finfo(np.float64).max


def _edges(cm):
    return np.ascontiguousarray(np.asarray(sorted(set(cm)), dtype=np.uint16).reshape(-1, 2))


def gpu_build_table(n, cm, w, exact):
    L = _abi.load()
    e = _edges(cm)
    w = np.ascontiguousarray(w, dtype=np.float64)
    out = np.empty((n, n))
    _abi.check(L, L.qmg_build_table(n, e.ctypes.data, len(e), w.ctypes.data, int(exact),
                                    out.ctypes.data, 0))
    return out


def gpu_single_skip(n, cm, d, rev):
    L = _abi.load()
    e = _edges(cm)
    d = np.ascontiguousarray(d, dtype=np.float64)
    out = np.empty((n, n))
    _abi.check(L, L.qmg_build_single_edge_skip_table(n, d.ctypes.data, e.ctypes.data, len(e),
                                                     float(rev), out.ctypes.data, 0))
    return out


def gpu_edge_skip(n, cm, w, exact):
    L = _abi.load()
    e = _edges(cm)
    w = np.ascontiguousarray(w, dtype=np.float64)
    out = np.zeros((n + 1, n, n))
    k = C.c_size_t(0)
    _abi.check(L, L.qmg_build_edge_skip_table(n, e.ctypes.data, len(e), w.ctypes.data, int(exact),
                                              out.ctypes.data, n + 1, C.byref(k), 0))
    return out[: k.value].copy()


random_weighted = parity.random_weighted


def _true_shortest(n, cm, w):
    d = np.full((n, n), np.inf)
    for a, b in cm:
        d[a, b] = min(d[a, b], w[a, b])
        d[b, a] = min(d[b, a], w[b, a])
    np.fill_diagonal(d, 0.0)
    for k in range(n):
        d = np.minimum(d, d[:, k:k + 1] + d[k:k + 1, :])
    return d


@pytest.mark.parametrize("n,extra,seed", [(4, 1, 0), (10, 5, 1), (33, 20, 2), (70, 60, 3), (129, 100, 4)])
def test_build_table_integer_weights(n, extra, seed):
    """K1' replays the reference's Dijkstra (bit for bit, any weights); K1
    (Floyd-Warshall) gives the true shortest paths, which is what the reference
    returns for uniform weights only."""
    cm, w = random_weighted(n, extra, seed, integer=True)
    assert np.array_equal(gpu_build_table(n, cm, w, exact=True), parity.ref().build_table(n, cm, w))
    assert np.array_equal(gpu_build_table(n, cm, w, exact=False), _true_shortest(n, cm, w))
    wu = np.where(w < DBL_MAX, 30.0, DBL_MAX)
    assert np.array_equal(gpu_build_table(n, cm, wu, exact=False), parity.ref().build_table(n, cm, wu))


@pytest.mark.parametrize("integer", [True, False])
def test_build_table_replay_240_seeds(integer):
    """Seeds 100..339 include the ones on which the reference's table is not
    the table of shortest paths (129, 134, 167, 175, 182, ...)."""
    R = parity.ref()
    for seed in range(100, 340):
        n = 12 if seed % 3 else 30
        cm, w = random_weighted(n, 10 if n == 12 else 25, seed, integer)
        assert np.array_equal(gpu_build_table(n, cm, w, exact=True), R.build_table(n, cm, w)), seed


@pytest.mark.parametrize("n,extra,seed", [(7, 3, 5), (40, 30, 6), (127, 40, 7)])
def test_build_table_float_weights(n, extra, seed):
    cm, w = random_weighted(n, extra, seed, integer=False)
    assert np.array_equal(gpu_build_table(n, cm, w, exact=True), parity.ref().build_table(n, cm, w))
    np.testing.assert_allclose(gpu_build_table(n, cm, w, exact=False), _true_shortest(n, cm, w),
                               rtol=1e-9)


@pytest.mark.parametrize("rev", [0.0, 1.0, 4.0])
def test_single_edge_skip(rev):
    cm, w = random_weighted(50, 30, 8, integer=True)
    d = parity.ref().build_table(50, cm, w)
    want = parity.ref().build_single_edge_skip_table(d, cm, rev)
    assert np.array_equal(gpu_single_skip(50, cm, d, rev), want)


@pytest.mark.parametrize("n,extra,seed,integer",
                         [(10, 4, 9, True), (24, 10, 10, False)] +
                         [(12, 8, s, bool(s & 1)) for s in range(120, 220, 5)])
def test_edge_skip_tables(n, extra, seed, integer):
    cm, w = random_weighted(n, extra, seed, integer)
    want = parity.ref().build_edge_skip_table(n, cm, w)
    got = gpu_edge_skip(n, cm, w, exact=True)
    assert got.shape == want.shape
    assert np.array_equal(got, want)


def _arch_cases():
    cases = [(os.path.basename(f), None) for f in sorted(glob.glob(os.path.join(parity.ARCHS, "*.arch")))]
    cases += [("heavy_hex_27", ar.heavy_hex_27), ("heavy_hex_127", ar.heavy_hex_127),
              ("grid_8x8", lambda: ar.grid(8, 8)), ("line_129", lambda: ar.line(129)),
              ("grid_16x16", lambda: ar.grid(16, 16)),
              ("line_12_directed", lambda: ar.line(12, False))]
    # mixed directions (30 / 34 SWAP costs): the reference's Dijkstra is replayed
    for seed in range(12):
        cases.append((f"mixed_{seed}", lambda seed=seed: _mixed(seed)))
    return cases


def _mixed(seed):
    n = 10 + 3 * seed
    cm, _ = random_weighted(n, n // 2 + seed, 500 + seed, True)
    return n, cm


@pytest.mark.parametrize("name,gen", _arch_cases())
def test_architecture_distance_tables(name, gen):
    if gen is None:
        a = ar.Architecture().load_coupling_map(os.path.join(parity.ARCHS, name))
    else:
        n, cm = gen()
        a = ar.Architecture(n, cm)
    ra = parity.ref().arch_create(a.n_qubits, a.coupling_map)
    for inc in (False, True):
        assert np.array_equal(a.distance_table(inc), ra.distance_table(inc)), (name, inc)
    info = a.info()
    assert bool(info.bidirectional) == bool(ra.info.bidirectional)
    assert bool(info.unidirectional) == bool(ra.info.unidirectional)


def test_fidelity_tables_london():
    a = ar.Architecture().load_coupling_map(os.path.join(parity.ARCHS, "ibmq_london.arch"))
    a.load_properties(os.path.join(parity.ARCHS, "ibmq_london.csv"))
    ra = parity.ref().arch_load(os.path.join(parity.ARCHS, "ibmq_london.arch"),
                                os.path.join(parity.ARCHS, "ibmq_london.csv"))
    want = ra.fidelity_distance_tables()
    got = a.fidelity_distance_tables()
    assert got.shape == want.shape
    assert np.array_equal(got, want)  # stronger than the 1e-9 the spec asks for


def test_fidelity_tables_heavy_hex_127(tmp_path):
    n, cm = ar.heavy_hex_127()
    a = ar.Architecture(n, cm)
    csv = ar.synthetic_calibration_csv(n, cm, 127)
    (tmp_path / "hh.arch").write_text(a.to_arch_text())
    (tmp_path / "hh.csv").write_text(csv)
    a.load_properties(str(tmp_path / "hh.csv"))
    ra = parity.ref().arch_load(str(tmp_path / "hh.arch"), str(tmp_path / "hh.csv"))
    want = ra.fidelity_distance_tables()
    got = a.fidelity_distance_tables()
    assert got.shape == want.shape
    np.testing.assert_allclose(got, want, rtol=1e-9, atol=0)
    assert np.array_equal(got, want)


def test_architecture_distance_tables_grid_32x32():
    """config 5's architecture (n = 1024); the reference needs about a minute."""
    n, cm = ar.grid(32, 32)
    a = ar.Architecture(n, cm)
    ra = parity.ref_arch_cached(n, cm)
    assert np.array_equal(a.distance_table(False), ra.distance_table(False))
    assert np.array_equal(a.distance_table(True), ra.distance_table(True))


@pytest.mark.parametrize("seed", range(40))
def test_fidelity_tables_random_calibration_grid_4x4(tmp_path, seed):
    """heterogeneous calibration: SWAP costs differ per edge, so the k-edge-skip
    tables start from the replayed Dijkstra (1 of 40 such tables
    differed with a true shortest-path kernel)."""
    n, cm = ar.grid(4, 4)
    a = ar.Architecture(n, cm)
    (tmp_path / "g.arch").write_text(a.to_arch_text())
    (tmp_path / "g.csv").write_text(ar.synthetic_calibration_csv(n, cm, 9000 + seed))
    a.load_properties(str(tmp_path / "g.csv"))
    ra = parity.ref().arch_load(str(tmp_path / "g.arch"), str(tmp_path / "g.csv"))
    want = ra.fidelity_distance_tables()
    got = a.fidelity_distance_tables()
    assert got.shape == want.shape
    assert np.array_equal(got, want)


# ---- the reference's own Architecture tests (test/sc/test_architecture.cpp:274-960)
def _arch_goldens():
    import json
    with open(os.path.join(parity.ROOT, "tests", "golden", "architecture_tests.json")) as f:
        return json.load(f)["tests"]


@pytest.mark.parametrize("g", _arch_goldens(), ids=lambda g: g["test"])
def test_reference_architecture_tests_on_the_device(g):
    from test_oracle import check_architecture_golden
    props = ar.Properties()
    for q, r in enumerate(g["single_qubit_error"]):
        props.set_single_qubit_error_rate(q, "x", r)
    for a_, b_, r in g["two_qubit_error"]:
        props.set_two_qubit_error_rate(a_, b_, r)
    a = ar.Architecture(g["n"], [tuple(e) for e in g["coupling_map"]], props)
    check_architecture_golden(g, a.fidelity_distance_tables(), a.swap_fidelity_costs())
    # and bit for bit what the compiled reference builds
    ra = parity.ref().arch_create(g["n"], [tuple(e) for e in g["coupling_map"]])
    rate = {(x, y): r for x, y, r in g["two_qubit_error"]}
    ra.set_error_rates(g["single_qubit_error"], [rate[tuple(int(v) for v in e)] for e in ra.edges])
    assert np.array_equal(a.fidelity_distance_tables(), ra.fidelity_distance_tables())


def test_distance_cheapest_path_on_the_device():
    """test_architecture.cpp:928-960; a mixed 30 / 34 architecture, so the table
    comes from the replay of the reference's Dijkstra (K1')."""
    n, cm, nr = parity.cheapest_path_architecture()
    a = ar.Architecture(n, cm)
    d = a.distance_table(True)
    assert d.shape == (n, n) and abs(d[0][nr] - nr * 30.0) <= 1e-6
    ra = parity.ref().arch_create(n, cm)
    assert np.array_equal(d, ra.distance_table(True))
    assert np.array_equal(a.distance_table(False), ra.distance_table(False))
