"""Runs the host debug build of the kernel source under ASan + UBSan over a
fifth of the golden cases (compute-sanitizer is not available on the GPU pool):

    ASAN=$(g++ -print-file-name=libasan.so); UBSAN=$(g++ -print-file-name=libubsan.so)
    LD_PRELOAD="$ASAN:$UBSAN" ASAN_OPTIONS=detect_leaks=0 python tests/emu/asan_check.py
"""
import sys, os
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE)); sys.path.insert(0, HERE)
import emu_backend
emu_backend.build(asan=True)
import ctypes as C
emu_backend._lib = None
orig_build = emu_backend.build
emu_backend.build = lambda asan=False: orig_build(asan=True)
import parity, test_emu_parity as T
bad = 0
for g in T.RUNS[::5]:
    be = T.emu_for(g.meta["arch"], g.meta["calibration"])
    msg, _ = parity.run_golden(be, be.lib, be.n, g)
    bad += bool(msg)
print("asan/ubsan run over", len(T.RUNS[::5]), "golden cases, mismatches", bad)
