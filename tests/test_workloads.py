"""CPU suite: the synthetic workloads of configs 3 and 4 (shared by bench.py
and the GPU parity tests) and the option / statistics structs of the ABI."""
import ctypes as C
import os
import re

import numpy as np

import parity
from qmap_b200 import _abi, workloads
from qmap_b200 import architecture as ar
from qmap_b200 import circuit as qcirc


def test_c4_circuits_are_seeded_and_distinct():
    a = workloads.c4_circuit(5, 30)
    b = workloads.c4_circuit(5, 30)
    c = workloads.c4_circuit(6, 30)
    assert a.n_qubits == 127
    assert [o.qubits for o in a.ops] == [o.qubits for o in b.ops]
    assert [o.qubits for o in a.ops] != [o.qubits for o in c.ops]
    assert sum(1 for o in a.ops if len(o.qubits) == 2) == 30
    # a longer instance extends a shorter one of the same seed? (no: the draw is
    # vectorised per length) - instances are identified by (seed, n_cx)
    ctl, tgt, identity = a.mapper_view(127)
    assert len(ctl) == len(a.ops) and identity.shape == (127,)


def test_c4_files_round_trip(tmp_path):
    arch_path, csv_path = workloads.c4_files(str(tmp_path))
    n, cm = ar.heavy_hex_127()
    tok = open(arch_path).read().split()
    assert int(tok[0]) == n == 127
    edges = {(int(tok[i]), int(tok[i + 1])) for i in range(1, len(tok), 2)}
    assert edges == set(cm)
    assert open(csv_path).read() == ar.synthetic_calibration_csv(n, cm, workloads.C4_CALIBRATION_SEED)


def test_c3_instances_are_placements_of_the_suite():
    suite = workloads.c3_suite(parity.CIRCUITS)
    assert len(suite) >= 20
    k = len(suite)
    seen = set()
    for seed in (0, 1, k, k + 1, 5 * k + 3):
        name, circ = workloads.c3_instance(parity.CIRCUITS, seed)
        base_name, base = suite[seed % k]
        assert name == base_name and circ.n_qubits == 27
        assert len(circ.ops) == len(base.ops)
        assert all(0 <= q < 27 for o in circ.ops for q in o.qubits)
        assert len(circ.used_qubits()) == len(base.used_qubits())
        # same gate structure up to the relabelling
        perm = {}
        for o, p in zip(base.ops, circ.ops):
            for q, r in zip(o.qubits, p.qubits):
                assert perm.setdefault(q, r) == r
        assert len(set(perm.values())) == len(perm)
        seen.add(tuple(sorted(perm.items())) + (name,))
        again = workloads.c3_instance(parity.CIRCUITS, seed)[1]
        assert [o.qubits for o in again.ops] == [o.qubits for o in circ.ops]
    assert len(seen) == 5
    batch = workloads.c3_instances(parity.CIRCUITS, 2)
    assert len(batch) == 2 * k


def _c_struct_fields(name):
    """field names of `typedef struct { ... } name;` in include/qmap_b200.h"""
    src = open(os.path.join(parity.ROOT, "include", "qmap_b200.h")).read()
    body = re.search(r"typedef struct \{((?:(?!typedef struct).)*?)\} %s;" % name, src, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    return [re.split(r"[\s\*]+", d.strip())[-1] for d in body.split(";") if d.strip()]


def test_option_and_stats_structs_mirror_the_header():
    assert _c_struct_fields("qmg_route_options") == [f[0] for f in _abi.QmgRouteOptions._fields_]
    assert _c_struct_fields("qmg_route_stats") == [f[0] for f in _abi.QmgRouteStats._fields_]
    assert C.sizeof(_abi.QmgRouteOptions) == 8 * 4
    assert C.sizeof(_abi.QmgRouteStats) == 72


def test_bench_workloads_shard_disjointly():
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench", os.path.join(parity.ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    assert set(bench.WORKLOADS) == {"c2", "c3", "c4"}
    for world in (1, 2, 8):
        seeds = [set(bench.shard_seeds(r, world, 100)) for r in range(world)]
        assert all(len(s) == 100 for s in seeds)
        assert len(set().union(*seeds)) == 100 * world
