"""ctypes bindings of the CPU restatement (oracle/qmap_oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
CPU legs of bench.py; never by the product package.  The restatement speaks
the C ABI's structs (include/qmap_b200.h), so `OracleBackend` can stand in for
the CUDA backend of qmap_b200.routing.route_batch in the parity tests.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_LIB = os.path.join(HERE, "libqmaporacle.so")


def build() -> str:
    subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])
    return ORACLE_LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORACLE_LIB):
            build()
        L = C.CDLL(ORACLE_LIB)
        vp = C.c_void_p
        L.qo_arch_create.argtypes = [C.c_uint16, vp, C.c_size_t]
        L.qo_arch_create.restype = vp
        L.qo_arch_destroy.argtypes = [vp]
        L.qo_arch_destroy.restype = None
        L.qo_arch_set_fidelity.argtypes = [vp, vp, vp, vp]
        L.qo_arch_set_fidelity.restype = None
        L.qo_arch_tables.argtypes = [vp, vp, vp]
        L.qo_arch_tables.restype = None
        L.qo_arch_flags.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.qo_arch_flags.restype = None
        L.qo_arch_fidelity_levels.argtypes = [vp]
        L.qo_arch_fidelity_levels.restype = C.c_uint32
        L.qo_arch_fidelity_tables.argtypes = [vp, vp]
        L.qo_arch_fidelity_tables.restype = None
        L.qo_build_table.argtypes = [C.c_uint16, vp, C.c_size_t, vp, vp]
        L.qo_build_table.restype = None
        L.qo_single_edge_skip.argtypes = [C.c_uint16, vp, C.c_size_t, vp, C.c_double, vp]
        L.qo_single_edge_skip.restype = None
        L.qo_edge_skip_levels.argtypes = [C.c_uint16, vp, C.c_size_t, vp, C.c_uint32, vp]
        L.qo_edge_skip_levels.restype = C.c_uint32
        L.qo_route.argtypes = [vp, vp, vp, vp, C.c_uint32]
        L.qo_route.restype = C.c_int
        L.qo_version.restype = C.c_char_p
        _lib = L
    return _lib


def _edges(edges) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(edges, dtype=np.uint16).reshape(-1, 2))


def _f64(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float64)


def build_table(n, edges, weights) -> np.ndarray:
    """Dijkstra::buildTable (src/sc/utils.cpp:26-54)."""
    e, w = _edges(edges), _f64(weights)
    out = np.zeros((n, n))
    lib().qo_build_table(n, e.ctypes.data, len(e), w.ctypes.data, out.ctypes.data)
    return out


def build_single_edge_skip_table(dist, edges, reversal_cost) -> np.ndarray:
    """Dijkstra::buildSingleEdgeSkipTable (utils.cpp:147-183)."""
    d, e = _f64(dist), _edges(edges)
    n = d.shape[0]
    out = np.zeros((n, n))
    lib().qo_single_edge_skip(n, e.ctypes.data, len(e), d.ctypes.data, float(reversal_cost),
                              out.ctypes.data)
    return out


def build_edge_skip_table(n, edges, weights) -> np.ndarray:
    """Dijkstra::buildEdgeSkipTable (utils.cpp:91-145) -> (levels, n, n)."""
    e, w = _edges(edges), _f64(weights)
    out = np.zeros((n + 1, n, n))
    k = lib().qo_edge_skip_levels(n, e.ctypes.data, len(e), w.ctypes.data, n + 1, out.ctypes.data)
    return out[:k].copy()


class OracleArch:
    """Architecture(nQ, couplingMap) of the restatement."""

    def __init__(self, n, edges):
        self.n = int(n)
        self.edges = _edges(edges)
        self.h = C.c_void_p(lib().qo_arch_create(self.n, self.edges.ctypes.data, len(self.edges)))
        self.fidelity = False

    def __del__(self):
        if getattr(self, "h", None):
            lib().qo_arch_destroy(self.h)
            self.h = None

    def set_fidelity_costs(self, c1, c2, cswap):
        a, b, c = _f64(c1), _f64(c2), _f64(cswap)
        lib().qo_arch_set_fidelity(self.h, a.ctypes.data, b.ctypes.data, c.ctypes.data)
        self.fidelity = True

    def distance_table(self, include_reversal_cost=True) -> np.ndarray:
        d = np.zeros((self.n, self.n))
        dr = np.zeros((self.n, self.n))
        lib().qo_arch_tables(self.h, d.ctypes.data, dr.ctypes.data)
        return dr if include_reversal_cost else d

    def flags(self):
        b, u = C.c_int(), C.c_int()
        lib().qo_arch_flags(self.h, C.byref(b), C.byref(u))
        return bool(b.value), bool(u.value)

    def fidelity_distance_tables(self) -> np.ndarray:
        k = lib().qo_arch_fidelity_levels(self.h)
        out = np.zeros((k, self.n, self.n))
        lib().qo_arch_fidelity_tables(self.h, out.ctypes.data)
        return out


class OracleBackend:
    """Backend protocol of qmap_b200.routing.route_batch on the restatement:
    every instance of the batch is routed by qo_route into the caller's
    qmg_route_result buffers."""

    def __init__(self, arch: OracleArch):
        self.arch = arch
        self.n = arch.n

    def route(self, cfg, b, options=None):
        L = lib()
        for i in range(len(b.items)):
            cap = b.layer_cap[i]
            L.qo_route(self.arch.h, C.byref(cfg), C.byref(b.inst[i]), C.byref(b.res[i]), cap)
