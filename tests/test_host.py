"""CPU suite: host-side logic of the product (no compute calls) - the C-ABI
library loads and exports every symbol of include/qmap_b200.h, it refuses to
compute without a CUDA device (no CPU fallback), the layering / splitting
entry points, the OpenQASM-2 reader, the architecture / calibration parsers,
and the sharding of instances over ranks (gloo, world size 2)."""
import os
import re
import subprocess
import sys
import textwrap

import ctypes as C

import numpy as np
import pytest

import parity
from qmap_b200 import _abi, routing
from qmap_b200 import architecture as ar
from qmap_b200 import circuit as qcirc

ROOT = parity.ROOT


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "qmap_b200.h")).read()
    declared = set(re.findall(r"\b(qmg_[a-z0-9_]+)\s*\(", hdr))
    # type names / macros that look like calls are not functions
    declared = {d for d in declared if not d.endswith("_t")}
    assert len(declared) >= 20
    L = C.CDLL(_abi.LIB_PATH)
    missing = [d for d in sorted(declared) if not hasattr(L, d)]
    assert not missing, missing
    assert set(_abi.EXPORTS) <= declared | {"qmg_config_default"}


def test_no_cpu_fallback():
    """Without a CUDA device every compute entry point returns QMG_NO_DEVICE."""
    L = _abi.load()
    if L.qmg_device_count() > 0:
        pytest.skip("a CUDA device is present")
    h = C.c_void_p()
    e = np.array([[0, 1], [1, 0]], dtype=np.uint16)
    rc = L.qmg_arch_create(2, e.ctypes.data, 2, 0, C.byref(h))
    assert rc == _abi.QMG_NO_DEVICE
    assert b"no CPU fallback" in L.qmg_last_error()
    w = np.zeros((2, 2))
    out = np.zeros((2, 2))
    rc = L.qmg_build_table(2, e.ctypes.data, 2, w.ctypes.data, 0, out.ctypes.data, 0)
    assert rc == _abi.QMG_NO_DEVICE
    with pytest.raises(_abi.QmapError):
        ar.Architecture(2, [(0, 1), (1, 0)]).handle


def test_config_defaults_are_the_references():
    """include/sc/configuration/Configuration.hpp:30-119"""
    L = _abi.load()
    c = _abi.QmgConfig()
    L.qmg_config_default(C.byref(c))
    py = routing.make_config()
    for name, _ in _abi.QmgConfig._fields_:
        assert getattr(c, name) == getattr(py, name), name
    assert (c.heuristic, c.lookahead_heuristic, c.nr_lookaheads) == (0, 1, 15)
    assert (c.first_lookahead_factor, c.lookahead_factor) == (0.75, 0.5)
    assert (c.automatic_layer_splits, c.automatic_layer_splits_node_limit) == (1, 5000)
    assert (c.layering, c.initial_layout) == (0, 2)


def _layers_of(lc):
    lay = lc.layers
    nl = lay.n_layers
    off2 = np.ctypeslib.as_array(C.cast(lay.gate2q_offsets, C.POINTER(C.c_uint32)), (nl + 1,))
    g2 = np.ctypeslib.as_array(C.cast(lay.gates2q, C.POINTER(C.c_uint16)), (max(int(off2[-1]), 1), 4))
    return nl, off2.copy(), g2[: off2[-1]].copy()


def test_layering_individual_and_disjoint():
    """Mapper::createLayers (src/sc/Mapper.cpp:138-281)"""
    L = _abi.load()
    # cx(0,1) h(2) cx(1,0) cx(2,3) cx(0,1)
    c = np.array([0, -1, 1, 2, 0], dtype=np.int32)
    t = np.array([1, 2, 0, 3, 1], dtype=np.int32)
    ind = routing.LayeredCircuit(L, 4, c, t, 0)
    assert ind.n_layers == 5
    dq = routing.LayeredCircuit(L, 4, c, t, 1)
    nl, off2, g2 = _layers_of(dq)
    # disjoint qubits: {cx01, h2}, {cx10, cx23}, {cx01}
    assert nl == 3
    assert list(off2) == [0, 1, 3, 4]
    assert g2[0].tolist() == [0, 1, 1, 0]
    assert g2[1].tolist() == [0, 1, 0, 1] and g2[2].tolist() == [2, 3, 1, 0]
    blk = routing.LayeredCircuit(L, 4, c, t, 4)
    nl, off2, g2 = _layers_of(blk)
    # disjoint 2q blocks: the three gates on (0,1) share one block with (2,3)
    assert nl == 1
    assert g2.tolist() == [[0, 1, 2, 1], [2, 3, 1, 0]]
    assert blk.splittable(0) and not ind.splittable(0)
    blk.split(0)
    assert blk.n_layers == 2


def test_router_rejects_unsuitable_settings():
    """HeuristicMapper::checkParameters :149-167 (checkConfig in host_api.hpp,
    called by every routing entry point before any work)"""
    import emu_backend
    n, e = parity.read_arch_file("ibm_qx4")
    t = parity.golden_tables()
    be = emu_backend.EmuBackend(n, e, t["ibm_qx4_dist"], t["ibm_qx4_dist_rev"])
    c = np.array([0], dtype=np.int32)
    tt = np.array([1], dtype=np.int32)
    for kw in (dict(layering="odd_gates"), dict(layering="qubit_triangle"),
               dict(heuristic="fidelity_best_location", lookahead_heuristic="none")):
        cfg = routing.make_config(**kw)
        item = routing.RouteItem(routing.LayeredCircuit(be.lib, n, c, tt, cfg.layering))
        with pytest.raises(_abi.QmapError):
            routing.route_batch(be, n, cfg, [item])


def test_qasm_reader_and_mapper_view():
    text = textwrap.dedent("""\
        OPENQASM 2.0;
        include "qelib1.inc";
        qreg q[5];
        creg c[5];
        h q[0];
        cx q[4],q[0];
        rz(pi/4) q[2];
        cx q[0], q[2];
        measure q[0] -> c[0];
        """)
    circ = qcirc.parse_qasm(text)
    assert circ.n_qubits == 5 and len(circ.ops) == 4 and circ.measurements == [(0, 0)]
    assert circ.used_qubits() == [0, 2, 4]
    c, t, ident = circ.mapper_view(5)
    # idle qubits 1 and 3 are stripped: q2 -> logical 1, q4 -> logical 2;
    # the identity layout keeps the original indices of the used qubits
    assert c.tolist() == [-1, 2, -1, 0] and t.tolist() == [0, 0, 1, 1]
    assert ident.tolist() == [0, -1, 2, -1, 4]
    again = qcirc.parse_qasm(circ.to_qasm())
    assert [o.qubits for o in again.ops] == [o.qubits for o in circ.ops]


def test_arch_and_calibration_parsers_match_reference_costs():
    """Architecture::loadCouplingMap / loadProperties / createFidelityTable
    (src/sc/Architecture.cpp:57-81,108-166,232-294): the host-side -log2 costs
    equal the reference's bit for bit (tests/golden/tables_golden.npz)."""
    a = ar.Architecture().load_coupling_map(os.path.join(parity.ARCHS, "ibmq_london.arch"))
    a.load_properties(os.path.join(parity.ARCHS, "ibmq_london.csv"))
    t = parity.golden_tables()
    c1, c2, cs = (a.single_qubit_fidelity_costs(), a.two_qubit_fidelity_costs(),
                  a.swap_fidelity_costs())
    assert np.array_equal(c1, t["ibmq_london_c1"])
    assert np.array_equal(c2, t["ibmq_london_c2"])
    assert np.array_equal(cs, t["ibmq_london_cswap"])
    assert a.n_qubits == 5 and len(a.edges) == 8
    assert a.to_arch_text().split()[0] == "5"


def test_generated_architectures():
    n, cm = ar.heavy_hex_27()
    assert n == 27 and len(cm) == 56
    n, cm = ar.heavy_hex_127()
    assert n == 127 and len(cm) == 288
    n, cm = ar.grid(32, 32)
    assert n == 1024 and len(cm) == 3968


def test_pack_and_unpack_roundtrip():
    L = _abi.load()
    c = np.array([0, 1], dtype=np.int32)
    t = np.array([1, 2], dtype=np.int32)
    lc = routing.LayeredCircuit(L, 3, c, t, 0)
    b = routing.pack(3, [routing.RouteItem(lc, np.array([0, 1, 2], dtype=np.int16))], lean=True)
    assert b.res[0].layers is None or not b.res[0].layers
    assert b.res[0].layer_n_swaps
    outs = routing.unpack(b)
    assert outs[0].layers is None and len(outs[0].layer_n_swaps) == 2


SHARD_SCRIPT = """
import os, sys
sys.path.insert(0, {root!r})
import torch.distributed as dist
import bench
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
seeds = list(bench.shard_seeds(rank, world, 10))
got = [None] * world
dist.all_gather_object(got, seeds)
if rank == 0:
    flat = [s for part in got for s in part]
    assert sorted(flat) == list(range(10 * world)), flat
    assert all(len(p) == 10 for p in got)
    print("SHARDS_OK", flat[:3])
dist.destroy_process_group()
"""


def test_instance_sharding_world_size_2(tmp_path):
    """bench.py --gpus N: rank r routes seeds [r*B, (r+1)*B) - disjoint, equal
    sized, no data-path collective (weak scaling)."""
    script = tmp_path / "shard.py"
    script.write_text(SHARD_SCRIPT.format(root=ROOT))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29631")
    out = subprocess.run(
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
         "--master-addr", "127.0.0.1", "--master-port", "29631", str(script)],
        capture_output=True, text=True, env=env, timeout=600)
    assert "SHARDS_OK" in out.stdout, out.stdout + out.stderr


def _layers_arrays(lay):
    nl = int(lay.n_layers)
    o2 = np.ctypeslib.as_array(C.cast(lay.gate2q_offsets, C.POINTER(C.c_uint32)), (nl + 1,)).copy()
    o1 = np.ctypeslib.as_array(C.cast(lay.gate1q_offsets, C.POINTER(C.c_uint32)), (nl + 1,)).copy()
    g2 = (np.ctypeslib.as_array(C.cast(lay.gates2q, C.POINTER(C.c_uint16)), (int(o2[-1]) * 4,)).copy()
          if o2[-1] else np.zeros(0, np.uint16))
    g1 = (np.ctypeslib.as_array(C.cast(lay.gates1q, C.POINTER(C.c_uint16)), (int(o1[-1]) * 2,)).copy()
          if o1[-1] else np.zeros(0, np.uint16))
    return nl, int(lay.n_logical), o2, o1, g2, g1


@pytest.mark.parametrize("layering", [0, 1, 4])
def test_circuit_batch_layers_equal_layered_circuit(layering):
    """qmg_circuit_batch_create (CSR-only createLayers, host threads) against
    qmg_circuit_create (the per-layer containers of the reference), array by
    array, on random circuits and the golden suite."""
    L = _abi.load()
    rng = np.random.default_rng(layering)
    lists = []
    for k in range(60):
        n = int(rng.integers(2, 12))
        m = int(rng.integers(0, 80))
        c = np.full(m, -1, dtype=np.int32)
        t = rng.integers(0, n, m).astype(np.int32)
        two = rng.random(m) < 0.6
        for i in np.nonzero(two)[0]:
            c[i] = (t[i] + 1 + rng.integers(0, n - 1)) % n
        lists.append((c, t))
    suite = os.path.join(ROOT, "tests", "golden", "circuits")
    for f in sorted(x for x in os.listdir(suite) if x.endswith(".qasm"))[:8]:
        lists.append(qcirc.load_qasm(os.path.join(suite, f)).gate_arrays())
    batch = routing.CircuitBatch.from_gate_lists(L, 16, lists, layering)
    assert len(batch) == len(lists)
    for i, (c, t) in enumerate(lists):
        one = routing.LayeredCircuit(L, 16, c, t, layering)
        a, b = _layers_arrays(batch[i].layers), _layers_arrays(one.layers)
        assert a[0] == b[0] and a[1] == b[1], i
        for x, y in zip(a[2:], b[2:]):
            assert np.array_equal(x, y), i
        for x, y in zip(batch[i].layer_gates(), one.layer_gates()):
            assert np.array_equal(x, y), i


def test_circuit_batch_rejects_bad_operands():
    L = _abi.load()
    with pytest.raises(_abi.QmapError):
        routing.CircuitBatch.from_gate_lists(L, 4, [(np.array([0]), np.array([4]))], 0)
    with pytest.raises(_abi.QmapError):
        routing.CircuitBatch.from_gate_lists(L, 4, [(np.array([0]), np.array([1]))], 2)


def test_minimum_number_of_swaps_and_coupling_limit_match_the_reference():
    """Architecture::minimumNumberOfSwaps (Architecture.cpp:296-385)
    and getCouplingLimit (:480-569), host utilities, against the compiled
    reference on the .arch fixtures and random permutations / subsets."""
    rng = np.random.default_rng(4)
    for name in ("ibm_qx4.arch", "ibmq_london.arch", "ibmq_yorktown.arch", "ibmq_bogota.arch",
                 "ibmq_casablanca.arch"):
        path = os.path.join(parity.ARCHS, name)
        a = ar.Architecture().load_coupling_map(path)
        ra = parity.ref().arch_load(path)
        n = a.n_qubits
        assert a.get_coupling_limit() == ra.coupling_limit(), name
        for _ in range(12):
            k = int(rng.integers(2, n + 1))
            subset = sorted(int(x) for x in rng.choice(n, size=k, replace=False))
            assert a.get_coupling_limit(subset) == ra.coupling_limit(subset), (name, subset)
            perm = [subset[i] for i in rng.permutation(k)]
            # (the reference's search only ends when the goal can be reached:
            # compare where the chosen qubits are connected among themselves)
            if a.get_coupling_limit(subset) > 0 and _connected(a, subset):
                want = ra.minimum_number_of_swaps(perm)
                assert a.minimum_number_of_swaps(perm) == want, (name, perm)
                if want > 1:
                    assert a.minimum_number_of_swaps(perm, want - 1) == ra.minimum_number_of_swaps(perm, want - 1)
    for n_, cm in (ar.heavy_hex_27(), ar.grid(4, 5), ar.line(9)):
        a = ar.Architecture(n_, cm)
        ra = parity.ref().arch_create(n_, sorted(cm))
        assert a.get_coupling_limit() == ra.coupling_limit()


def _connected(a, subset):
    s = set(subset)
    adj = {q: set() for q in s}
    for x, y in a.coupling_map:
        if x in s and y in s:
            adj[x].add(y)
            adj[y].add(x)
    seen, todo = {subset[0]}, [subset[0]]
    while todo:
        u = todo.pop()
        for v in adj[u]:
            if v not in seen:
                seen.add(v)
                todo.append(v)
    return seen == s


# ---------------------------------------------------------------------------
# the library's OpenQASM 2 reader (csrc/qasm_reader.hpp, qmg_qasm_*) against
# the Python reader (qmap_b200/circuit.py), which the golden runs pin to the
# reference's own import of the same files
# ---------------------------------------------------------------------------
def _qasm_files():
    import glob
    files = sorted(glob.glob(os.path.join(parity.CIRCUITS, "*.qasm")))
    full = os.path.join(os.path.dirname(parity.CIRCUITS), "circuits_full")
    big = sorted(glob.glob(os.path.join(full, "*.qasm")), key=os.path.getsize)
    return files + big[:60] + big[-3:]


def test_native_qasm_reader_equals_the_python_reader():
    from qmap_b200 import circuit as qc
    L = _abi.load()
    n_gates = 0
    for path in _qasm_files():
        want = qc.load_qasm(path)
        got = qc.load_qasm_native(L, path)
        assert (got.n_qubits, got.n_clbits, got.n_gates) == (want.n_qubits, want.n_clbits, len(want.ops)), path
        wc, wt = want.gate_arrays()
        c, t, k, p, npar, names = got.gates()
        assert np.array_equal(c, wc) and np.array_equal(t, wt), path
        assert [names[i] for i in k] == [o.name for o in want.ops], path
        for i in np.nonzero(npar)[0]:
            assert tuple(p[i, :npar[i]]) == tuple(want.ops[i].params), (path, i)   # bit for bit
        assert all(len(o.params) == npar[i] for i, o in enumerate(want.ops)), path
        for n_phys in (want.n_qubits, 27):
            a, b = got.mapper_view(n_phys), want.mapper_view(n_phys)
            assert all(np.array_equal(x, y) for x, y in zip(a, b)), path
        n_gates += got.n_gates
    assert n_gates > 3000  # (the committed subset alone has 3717 gates)


def test_native_qasm_reader_statements_and_errors():
    from qmap_b200 import circuit as qc
    L = _abi.load()
    text = """OPENQASM 2.0; include "qelib1.inc";  // header
    qreg a[2]; qreg b[3]; creg c[5]; creg d[1];
    h a[0]; CX a[1],b[2]; cx b[0] , a[0];
    u3(pi/2, -pi/4 , 0.25e1) b[1];  rz( (1+2)*pi/8 - 3/4 ) b[2]; u1(-(pi)) a[1]; u2(0,pi) a[0];
    barrier a[0],b[1];
    measure b -> c; measure a[1] -> d[0];
    t b[2]; // trailing comment
    """
    want = qc.parse_qasm(text)
    got = qc.NativeQasm(L, text=text)
    circ = got.to_circuit()
    assert (circ.n_qubits, circ.n_clbits) == (want.n_qubits, want.n_clbits) == (5, 6)
    assert [(o.name, o.qubits, o.params) for o in circ.ops] == [(o.name, o.qubits, o.params) for o in want.ops]
    assert circ.measurements == want.measurements == [(2, 0), (3, 1), (4, 2), (1, 5)]
    # qubit 4 = b[2] is used, everything is: identity on 5 of 8 physical qubits
    assert got.mapper_view(8)[2].tolist() == [0, 1, 2, 3, 4, -1, -1, -1]
    # an idle qubit in the middle moves the logical indices above it down (Mapper.cpp:44-53)
    idle = qc.NativeQasm(L, text="qreg q[4]; cx q[0],q[3]; h q[1];")
    c, t, ident = idle.mapper_view(5)
    assert (c.tolist(), t.tolist(), ident.tolist()) == ([0, -1], [2, 1], [0, 1, -1, 3, -1])
    w = qc.parse_qasm("qreg q[4]; cx q[0],q[3]; h q[1];").mapper_view(5)
    assert (w[0].tolist(), w[1].tolist(), w[2].tolist()) == ([0, -1], [2, 1], [0, 1, -1, 3, -1])
    for bad, msg in (("qreg q[3]; ccx q[0],q[1],q[2];", "more than one control or an unsupported gate \\(ccx\\)"),
                     ("qreg q[2]; cx q[0];", "bad cx"),
                     ("qreg q[2]; h q[0],q[1];", "bad 1q gate"),
                     ("qreg q[2]; h r[0];", "bad qubit argument"),
                     ("qreg q[2]; rz(pi//2) q[0];", "parse"),
                     ("qreg q[2]; rz(2*) q[0];", "parse")):
        with pytest.raises(_abi.QmapError, match=msg):
            qc.NativeQasm(L, text=bad)
        with pytest.raises((ValueError, IndexError)):
            qc.parse_qasm(bad)
    with pytest.raises(_abi.QmapError, match="SWAP gates"):
        qc.NativeQasm(L, text="qreg q[2]; swap q[0],q[1];").mapper_view(2)
    with pytest.raises(_abi.QmapError, match="cannot open"):
        qc.load_qasm_native(L, "/nonexistent/x.qasm")


def test_circuit_batch_from_qasm_files_equals_the_python_path():
    """Files -> CSR layers through the library alone (qmg_qasm_read_file,
    qmg_qasm_mapper_view, qmg_circuit_batch_create) against the Python reader
    + qmg_circuit_create."""
    L = _abi.load()
    paths = _qasm_files()[:30]
    n = 27
    for layering in (0, 1, 4):
        batch, idents = routing.CircuitBatch.from_qasm_files(L, n, paths, layering)
        assert len(batch) == len(paths)
        for i, p in enumerate(paths):
            c, t, ident = qcirc.load_qasm(p).mapper_view(n)
            assert np.array_equal(idents[i], ident), p
            one = routing.LayeredCircuit(L, n, c, t, layering)
            a, b = _layers_arrays(batch[i].layers), _layers_arrays(one.layers)
            assert a[0] == b[0] and a[1] == b[1], p
            for x, y in zip(a[2:], b[2:]):
                assert np.array_equal(x, y), p


def test_native_qasm_reader_survives_malformed_input():
    """The reader takes user text through the C ABI: whatever it is given, it
    answers with a status (mutations of a valid file, truncations, junk)."""
    from qmap_b200 import circuit as qc
    L = _abi.load()
    base = open(os.path.join(parity.CIRCUITS, "3_17_13.qasm")).read()
    rng = np.random.default_rng(7)
    junk = list("[](),;->.0123456789 piqregcxmeasurebarrier\n\t*/+-eE\"")
    n_ok = n_err = 0
    texts = [base[:k] for k in range(0, len(base), 7)]
    for _ in range(1500):
        t = list(base)
        for _ in range(int(rng.integers(1, 6))):
            op = int(rng.integers(0, 3))
            pos = int(rng.integers(0, len(t)))
            if op == 0:
                t[pos] = junk[int(rng.integers(0, len(junk)))]
            elif op == 1:
                del t[pos]
            else:
                t.insert(pos, junk[int(rng.integers(0, len(junk)))])
        texts.append("".join(t))
    texts += ["qreg q[3]; cx q[0],q[7];", "qreg q[2]; creg c[1]; measure q[1] -> c[5];",
              "qreg q[99999999999999999999];", "qreg q[2]; creg c[1]; measure q -> c;",
              "qreg q[20000000]; qreg r[20000000];", "qreg q[2]; rz(((((((1))))))) q[0];",
              "qreg q[2]; rz(1e999) q[0];", "qreg q[1]; u3(1,2,3,4) q[0];", ";;;;", "(", "qreg", "measure ->"]
    for text in texts:
        try:
            n = qc.NativeQasm(L, text=text)
            n.mapper_view(5)
            n.gates()
            n.measurements()
            n_ok += 1
        except _abi.QmapError:
            n_err += 1
    assert n_ok > 50 and n_err > 500


def test_c_example_compiles_against_the_header_and_library(tmp_path):
    """examples/route_qasm.c: include/qmap_b200.h is plain C11 (-pedantic), the
    example links against libqmap_b200.so and runs the host part of the path
    (QASM reader, mapper view, createLayers); without a GPU it stops at the
    device check with exit code 3 - there is no CPU fallback."""
    import shutil
    import subprocess
    import torch
    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None:
        pytest.skip("no C compiler")
    libdir = os.path.join(ROOT, "mqt-qmap_b200", "csrc")
    exe = str(tmp_path / "route_qasm")
    subprocess.check_call([cc, "-std=c11", "-Wall", "-Wextra", "-pedantic", "-Werror",
                           "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "examples", "route_qasm.c"),
                           "-L" + libdir, "-lqmap_b200", "-Wl,-rpath," + libdir, "-o", exe])
    r = subprocess.run([exe, os.path.join(parity.CIRCUITS, "3_17_13.qasm"),
                        os.path.join(parity.ARCHS, "ibm_qx4.arch")], capture_output=True, text=True, timeout=120)
    if torch.cuda.is_available():
        assert r.returncode == 0 and "SWAP" in r.stdout, r.stderr
    else:
        assert r.returncode == 3 and "no CUDA device" in r.stderr, (r.returncode, r.stderr)
    bad = subprocess.run([exe, os.path.join(parity.ARCHS, "ibm_qx4.arch"),
                          os.path.join(parity.ARCHS, "ibm_qx4.arch")], capture_output=True, text=True, timeout=120)
    assert bad.returncode == 1 and "reading the circuit" in bad.stderr
