"""-m gpu: the table kernels (K1 blocked Floyd-Warshall, K1' exact label
correcting, K2 single-edge-skip, K2' k-edge-skip) against the compiled
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


def random_weighted(n, extra, seed, integer):
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


@pytest.mark.parametrize("n,extra,seed", [(4, 1, 0), (10, 5, 1), (33, 20, 2), (70, 60, 3), (129, 100, 4)])
def test_build_table_integer_weights_both_kernels(n, extra, seed):
    cm, w = random_weighted(n, extra, seed, integer=True)
    want = parity.ref().build_table(n, cm, w)
    assert np.array_equal(gpu_build_table(n, cm, w, exact=False), want)
    assert np.array_equal(gpu_build_table(n, cm, w, exact=True), want)


@pytest.mark.parametrize("n,extra,seed", [(7, 3, 5), (40, 30, 6), (127, 40, 7)])
def test_build_table_float_weights(n, extra, seed):
    cm, w = random_weighted(n, extra, seed, integer=False)
    want = parity.ref().build_table(n, cm, w)
    # exact association: bit for bit; Floyd-Warshall: <= 1e-9 relative
    assert np.array_equal(gpu_build_table(n, cm, w, exact=True), want)
    np.testing.assert_allclose(gpu_build_table(n, cm, w, exact=False), want, rtol=1e-9)


@pytest.mark.parametrize("rev", [0.0, 1.0, 4.0])
def test_single_edge_skip(rev):
    cm, w = random_weighted(50, 30, 8, integer=True)
    d = parity.ref().build_table(50, cm, w)
    want = parity.ref().build_single_edge_skip_table(d, cm, rev)
    assert np.array_equal(gpu_single_skip(50, cm, d, rev), want)


@pytest.mark.parametrize("n,extra,seed,integer", [(10, 4, 9, True), (24, 10, 10, False)])
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
              ("line_12_directed", lambda: ar.line(12, False))]
    return cases


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
