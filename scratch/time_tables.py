"""Times the table kernels (qmg_arch_create) on the architectures of configs 2-5."""
import sys, os, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "mqt-qmap_b200"))
import numpy as np
from qmap_b200 import architecture as ar, _abi
out = {}
for name, (n, cm) in (("heavy_hex_27", ar.heavy_hex_27()), ("heavy_hex_127", ar.heavy_hex_127()),
                      ("grid_32x32", ar.grid(32, 32))):
    for rep in range(3):
        a = ar.Architecture(n, cm)
        t0 = time.perf_counter(); h = a.handle; t1 = time.perf_counter()
        info = a.info()
        ms = float(info.table_build_ms) if hasattr(info, "table_build_ms") else None
        d = a.distance_table(True)
        out[name] = {"n": n, "edges": len(cm), "create_wall_ms": 1e3 * (t1 - t0), "kernel_ms": ms,
                     "max_dist": float(d.max())}
        del a
    if name == "heavy_hex_127":
        a = ar.Architecture(n, cm)
        csv = ar.synthetic_calibration_csv(n, cm, 127)
        p = "/tmp/cal127.csv"; open(p, "w").write(csv)
        a.load_properties(p)
        t0 = time.perf_counter(); h = a.handle; f = a.fidelity_distance_tables(); t1 = time.perf_counter()
        out[name + "_fidelity"] = {"levels": int(f.shape[0]), "wall_ms": 1e3 * (t1 - t0),
                                   "kernel_ms": float(a.info().table_build_ms) if hasattr(a.info(), "table_build_ms") else None}
print(json.dumps(out))
