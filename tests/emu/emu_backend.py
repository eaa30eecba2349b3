"""Test-only backend: the host debug build of the routing kernel's source
(tests/emu/libqmg_emu.so).  Cost tables are taken from arrays the test
supplies (normally the compiled reference, oracle/_ref)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "mqt-qmap_b200"))

from qmap_b200 import _abi  # noqa: E402


def build(asan: bool = False) -> str:
    # QMG_EMU_BLK="4,8": open-list blocks of 4 entries from 8 open nodes on, so
    # that the small golden cases exercise the blocked open list
    blk = os.environ.get("QMG_EMU_BLK", "")
    # QMG_EMU_DEFS="QMG_FLIGHT=4,QMG_WIDE=4": other settings of the kernel's
    # tuning macros (they must not change any result)
    defs = [d for d in os.environ.get("QMG_EMU_DEFS", "").split(",") if d]
    tag = (("_blk_" + blk.replace(",", "_")) if blk else "") + \
        (("_" + "_".join(d.replace("=", "") for d in defs)) if defs else "")
    out = os.path.join(HERE, "libqmg_emu_asan.so" if asan else f"libqmg_emu{tag}.so")
    src = os.path.join(HERE, "emu.cpp")
    deps = [src] + [os.path.join(ROOT, "mqt-qmap_b200", "csrc", f)
                    for f in ("route_core.hpp", "layers.hpp", "host_api.hpp", "flatten.hpp",
                              "dijkstra_core.hpp", "flat_layers.hpp", "qasm_reader.hpp")]
    deps.append(os.path.join(ROOT, "include", "qmap_b200.h"))
    if os.path.exists(out) and all(os.path.getmtime(out) >= os.path.getmtime(d) for d in deps):
        return out
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    flags = ["-std=c++17", "-fPIC", "-shared", "-Wall", "-Wextra", "-g"]
    flags += ["-O1", "-fsanitize=address,undefined", "-fno-omit-frame-pointer"] if asan else ["-O2"]
    if blk:
        # "4,8,2": in addition the layout hash table from 3 open nodes on
        parts = blk.split(",")
        flags += [f"-DQMG_BLK={int(parts[0])}", f"-DQMG_BLK_MIN={int(parts[1])}"]
        if len(parts) > 2:
            flags += [f"-DQMG_SCAN_MAX={int(parts[2])}"]
    flags += ["-D" + d for d in defs]
    subprocess.check_call([cxx, *flags, "-o", out, src])
    return out


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        _abi.bind_host_api(L)
        L.qemu_arch_create.argtypes = [C.c_uint16, C.c_void_p, C.c_size_t, C.c_void_p,
                                       C.c_void_p, C.POINTER(C.c_void_p)]
        L.qemu_arch_set_fidelity.argtypes = [C.c_void_p] * 5 + [C.c_uint32]
        L.qemu_arch_destroy.argtypes = [C.c_void_p]
        L.qemu_arch_destroy.restype = None
        L.qemu_route_batch.argtypes = [C.c_void_p, C.POINTER(_abi.QmgConfig), C.c_size_t,
                                       C.c_void_p, C.c_void_p, C.c_uint32]
        L.qemu_build_table.argtypes = [C.c_uint16, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def build_table(n, edges, weights):
    """Dijkstra::buildTable by the device function of table kernel K1' (host build)."""
    L = lib()
    e = np.ascontiguousarray(edges, dtype=np.uint16).reshape(-1, 2)
    w = np.ascontiguousarray(weights, dtype=np.float64)
    out = np.empty((n, n), dtype=np.float64)
    _abi.check(L, L.qemu_build_table(n, e.ctypes.data, len(e), w.ctypes.data, out.ctypes.data))
    return out


class EmuBackend:
    def __init__(self, n, edges, dist, dist_rev, fidelity=None, pool_nodes=0):
        """fidelity = (c1, c2, cswap, fid_tables) or None"""
        self.lib = lib()
        self.n = int(n)
        e = np.ascontiguousarray(edges, dtype=np.uint16).reshape(-1, 2)
        d = np.ascontiguousarray(dist, dtype=np.float64)
        dr = np.ascontiguousarray(dist_rev, dtype=np.float64)
        h = C.c_void_p()
        _abi.check(self.lib, self.lib.qemu_arch_create(
            self.n, e.ctypes.data, len(e), d.ctypes.data, dr.ctypes.data, C.byref(h)))
        self.h = h
        self.pool_nodes = pool_nodes
        if fidelity is not None:
            c1, c2, cs, fid = (np.ascontiguousarray(x, dtype=np.float64) for x in fidelity)
            self.lib.qemu_arch_set_fidelity(h, c1.ctypes.data, c2.ctypes.data, cs.ctypes.data,
                                            fid.ctypes.data, fid.shape[0])

    @classmethod
    def from_ref_arch(cls, ra, pool_nodes=0):
        """ra: oracle.refapi.RefArch"""
        fid = None
        if ra.info.fidelity_available:
            fid = (ra.single_qubit_fidelity_costs(), ra.two_qubit_fidelity_costs(),
                   ra.swap_fidelity_costs(), ra.fidelity_distance_tables())
        return cls(ra.n, ra.edges, ra.distance_table(False), ra.distance_table(True),
                   fid, pool_nodes)

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.qemu_arch_destroy(self.h)
            self.h = None

    def route(self, cfg, b, options=None):
        rc = self.lib.qemu_route_batch(self.h, C.byref(cfg), len(b.items),
                                       C.addressof(b.inst), C.addressof(b.res),
                                       self.pool_nodes)
        _abi.check(self.lib, rc)
