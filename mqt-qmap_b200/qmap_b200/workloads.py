"""Synthetic workloads of BASELINE.json's configs 3 and 4, shared by bench.py
and the parity tests so that both route exactly the same instances.

config 3: the examples/*.qasm suite on a 27-qubit heavy-hex architecture,
          circuits x seeded relabellings (= different initial layouts)
config 4: seeded random 127-qubit circuits on a 127-qubit heavy-hex
          architecture with synthetic calibration data, fidelity-aware
          heuristic (fidelity_best_location, HeuristicMapper.cpp:1247-1354)
"""
from __future__ import annotations

import os
import tempfile

import numpy as np

from . import architecture as ar
from . import circuit as qcirc

C4_CALIBRATION_SEED = 127
C4_CONFIG = dict(heuristic="fidelity_best_location", lookahead_heuristic="none",
                 initial_layout="identity")
C3_CONFIG = dict(initial_layout="identity")


def c4_files(directory: str | None = None):
    """(arch file, calibration csv) of the config-4 architecture, written to
    `directory` (a fresh temporary directory by default)."""
    n, cm = ar.heavy_hex_127()
    d = directory or tempfile.mkdtemp(prefix="qmap_c4_")
    os.makedirs(d, exist_ok=True)
    arch_path = os.path.join(d, "heavy_hex_127.arch")
    csv_path = os.path.join(d, "heavy_hex_127.csv")
    with open(arch_path, "w") as f:
        f.write(f"{n}\n" + "\n".join(f"{a} {b}" for a, b in sorted(cm)) + "\n")
    with open(csv_path, "w") as f:
        f.write(ar.synthetic_calibration_csv(n, cm, C4_CALIBRATION_SEED))
    return arch_path, csv_path


def c4_circuit(seed: int, n_cx: int) -> qcirc.Circuit:
    """Instance `seed` of config 4: n_cx CX on uniformly random pairs of the
    127 qubits, 1q gates in between (circuit.random_circuit)."""
    return qcirc.random_circuit(127, n_cx, 1000 + seed)


_c3_bases = {}


def c3_suite(circuit_dir: str):
    """[(file name, Circuit)] of the suite, sorted by file name (cached)."""
    if circuit_dir not in _c3_bases:
        files = sorted(f for f in os.listdir(circuit_dir) if f.endswith(".qasm"))
        _c3_bases[circuit_dir] = [(f, qcirc.load_qasm(os.path.join(circuit_dir, f))) for f in files]
    return _c3_bases[circuit_dir]


def c3_instance(circuit_dir: str, seed: int, n_physical: int = 27):
    """Instance `seed` of config 3: circuit seed % |suite|, widened to
    n_physical qubits and relabelled by a seeded injection of its qubits into
    them - with the identity initial layout every seed starts the circuit from
    a different placement."""
    suite = c3_suite(circuit_dir)
    f, base = suite[seed % len(suite)]
    rng = np.random.default_rng([seed, len(f), base.n_qubits])
    wide = qcirc.Circuit(n_physical, ops=list(base.ops), name=base.name,
                         n_clbits=base.n_clbits, measurements=list(base.measurements))
    return f, wide.relabel(rng.permutation(n_physical))


def c3_passes(circuit_dir: str, seed: int) -> int:
    """Bidirectional pseudo-routing passes of instance `seed`: every placement
    of the suite is routed forward only (even rounds over the suite) and with
    one bidirectional pass (odd rounds), BASELINE.md section 3.3."""
    return (seed // len(c3_suite(circuit_dir))) % 2


def c3_instance_passes(circuit_dir: str, seed: int, n_physical: int = 27):
    f, circ = c3_instance(circuit_dir, seed, n_physical)
    return f, circ, c3_passes(circuit_dir, seed)


_c3_arrays = {}


def c3_arrays(circuit_dir: str):
    """[(file name, n_qubits, control, target)] of the suite as gate arrays,
    sorted by file name; parsed once and kept in <dir>/_gate_arrays.npz (the
    full suite is 2 M gates: 13 s of parsing)."""
    if circuit_dir in _c3_arrays:
        return _c3_arrays[circuit_dir]
    files = sorted(f for f in os.listdir(circuit_dir) if f.endswith(".qasm"))
    cache = os.path.join(circuit_dir, "_gate_arrays.npz")
    out = None
    if os.path.exists(cache):
        try:
            z = np.load(cache)
            if list(z["files"]) == files:
                off = z["offsets"]
                out = [(f, int(z["n_qubits"][i]), z["control"][off[i]:off[i + 1]], z["target"][off[i]:off[i + 1]])
                       for i, f in enumerate(files)]
        except Exception:  # noqa: BLE001 (a stale or broken cache is rebuilt)
            out = None
    if out is None:
        out = []
        for f in files:
            circ = qcirc.load_qasm(os.path.join(circuit_dir, f))
            c, t = circ.gate_arrays()
            out.append((f, circ.n_qubits, c, t))
        try:
            off = np.concatenate(([0], np.cumsum([len(x[2]) for x in out]))).astype(np.int64)
            np.savez(cache, files=np.array(files), n_qubits=np.array([x[1] for x in out]), offsets=off,
                     control=np.concatenate([x[2] for x in out]), target=np.concatenate([x[3] for x in out]))
        except OSError:
            pass
    _c3_arrays[circuit_dir] = out
    return out


def c3_instance_arrays(circuit_dir: str, seed: int, n_physical: int = 27):
    """c3_instance_passes as gate arrays: (n_physical, control, target, passes);
    equal to c3_instance(...)[1].gate_arrays() without building Op lists."""
    suite = c3_arrays(circuit_dir)
    f, nq, c, t = suite[seed % len(suite)]
    rng = np.random.default_rng([seed, len(f), nq])
    perm = rng.permutation(n_physical).astype(np.int32)
    c2 = np.where(c >= 0, perm[np.maximum(c, 0)], -1).astype(np.int32)
    return n_physical, c2, perm[t].astype(np.int32), (seed // len(suite)) % 2


def c3_directory(root: str) -> str:
    """The circuit suite of config 3: all of the reference's examples/*.qasm
    where tools/fetch_examples.py has copied them (tests/golden/circuits_full,
    git-ignored, travels to the GPU box), else the committed 22-circuit subset
    (tests/golden/circuits)."""
    full = os.path.join(root, "tests", "golden", "circuits_full")
    if os.path.isdir(full) and any(f.endswith(".qasm") for f in os.listdir(full)):
        return full
    return os.path.join(root, "tests", "golden", "circuits")


def c3_instances(circuit_dir: str, copies: int, first_seed: int = 0, n_physical: int = 27):
    """[(file name, Circuit)] for seeds first_seed .. first_seed + copies * |suite|."""
    k = len(c3_suite(circuit_dir))
    return [c3_instance(circuit_dir, s, n_physical)
            for s in range(first_seed, first_seed + copies * k)]
