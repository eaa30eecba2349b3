"""-m gpu, collected last: qmg_build_table_static (the table build that
allocates nothing: its scratch is a static device array inside the library)
against qmg_build_table's Floyd-Warshall path, through the C ABI, on calls in
a row that share the static scratch."""
import ctypes as C

import numpy as np
import pytest

import parity
from qmap_b200 import _abi

pytestmark = pytest.mark.gpu
MAX_N = 512  # QMG_STATIC_TABLE_MAX_QUBITS (include/qmap_b200.h)


def _edges(cm):
    return np.ascontiguousarray(np.asarray(sorted(set(cm)), dtype=np.uint16).reshape(-1, 2))


def _existing(L, n, e, w):
    out = np.empty((n, n))
    _abi.check(L, L.qmg_build_table(n, e.ctypes.data, len(e), w.ctypes.data, 0,
                                    out.ctypes.data, 0))
    return out


def _static(L, n, e, w):
    """Inputs and output are device buffers owned by the caller (torch here)."""
    import torch
    de = torch.from_numpy(e.view(np.int16).copy()).cuda()
    dw = torch.from_numpy(w).cuda()
    dout = torch.full((n, n), -7.0, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    _abi.check(L, L.qmg_build_table_static(n, de.data_ptr(), len(e), dw.data_ptr(),
                                           dout.data_ptr(), 0, None))
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return dout.cpu().numpy()


def test_static_table_build_matches_existing_entry_point_on_calls_in_a_row():
    L = _abi.load()
    # large, then small (stale cells of the larger call stay in the scratch),
    # then large again; integer and float weights; a disconnected pair
    cases = [(129, 100, 41, True), (7, 3, 42, False), (200, 150, 43, False), (33, 20, 44, True)]
    for n, extra, seed, integer in cases:
        cm, w = parity.random_weighted(n, extra, seed, integer)
        e = _edges(cm)
        want = _existing(L, n, e, w)
        got = _static(L, n, e, w)
        assert np.array_equal(got, want), (n, seed)   # bit for bit, as tests/test_gpu_tables.py
    # two components: unreachable pairs are -1 in both
    n = 10
    cm = [(0, 1), (1, 2), (5, 6), (6, 5)]
    w = np.full((n, n), parity.DBL_MAX)
    for a, b in cm:
        w[a, b] = w[b, a] = 30.0
    e = _edges(cm)
    got = _static(L, n, e, w)
    assert np.array_equal(got, _existing(L, n, e, w))
    assert got[0, 5] == -1.0 and got[0, 2] == 60.0


def test_static_table_build_rejects_sizes_beyond_the_scratch():
    L = _abi.load()
    import torch
    buf = torch.zeros(16, dtype=torch.float64, device="cuda")
    rc = L.qmg_build_table_static(MAX_N + 1, buf.data_ptr(), 0, buf.data_ptr(), buf.data_ptr(), 0, None)
    assert rc == _abi.QMG_BAD_ARG
    assert b"exceeds the static scratch" in L.qmg_last_error()
