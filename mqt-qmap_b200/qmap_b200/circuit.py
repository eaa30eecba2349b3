"""Minimal circuit IR for the heuristic-mapper path: an OpenQASM 2 reader for
the gate set of the reference's examples/*.qasm and a writer for mapped
circuits.  (The reference delegates this to mqt-core's QuantumComputation /
qasm3::Importer, which are outside the accelerated path.)
"""
from __future__ import annotations

import math
import os
import re
from dataclasses import dataclass, field

import numpy as np

ONE_QUBIT_GATES = {
    "h", "x", "y", "z", "s", "sdg", "t", "tdg", "id", "sx", "sxdg",
    "rx", "ry", "rz", "u1", "u2", "u3", "u", "p",
}
TWO_QUBIT_GATES = {"cx", "CX"}


@dataclass
class Op:
    name: str
    qubits: tuple  # (target,) or (control, target); for swap (q0, q1)
    params: tuple = ()

    @property
    def is_two_qubit(self) -> bool:
        return len(self.qubits) == 2


@dataclass
class Circuit:
    """Gate list over `n_qubits` qubits; measurements and barriers are kept
    apart (the mapper ignores them, src/sc/Mapper.cpp:146-149)."""
    n_qubits: int
    ops: list = field(default_factory=list)
    name: str = "circuit"
    n_clbits: int = 0
    measurements: list = field(default_factory=list)  # (qubit, clbit)

    def cx(self, c: int, t: int):
        self.ops.append(Op("cx", (int(c), int(t))))

    def gate(self, name: str, q: int, *params):
        self.ops.append(Op(name, (int(q),), tuple(params)))

    def h(self, q: int):
        self.gate("h", q)

    def used_qubits(self) -> list:
        """Qubits that are not idle: touched by a gate or measured."""
        used = set()
        for op in self.ops:
            used.update(op.qubits)
        used.update(q for q, _ in self.measurements)
        return sorted(used)

    def mapper_view(self, n_physical: int):
        """What the reference mapper makes of this circuit before routing:
        (control, target, identity_locations).

        Mapper's constructor strips idle qubits (src/sc/Mapper.cpp:44-53;
        mqt-core's stripIdleQubits(force) moves the logical indices above a
        removed qubit down by one) and createLayers reads every gate operand
        through `initialLayout` (Mapper.cpp:166-172), so operand q becomes the
        rank of q among the used qubits.  InitialLayout::Identity then places
        logical i on physical i for every *original* used index i
        (HeuristicMapper.cpp:176-183) - with an idle qubit in the middle that
        leaves the highest logical qubit(s) unplaced until their first gate,
        which the golden layouts of 4gt11_84 in the reference's
        test_heuristic.cpp pin."""
        if any(op.name == "swap" for op in self.ops):
            raise ValueError("Circuit contains SWAP gates. Please make sure that the input "
                             "circuit's gates are decomposed to the appropriate gate set!")
        used = self.used_qubits()
        rank = {q: i for i, q in enumerate(used)}
        c, t = self.gate_arrays()
        lut = np.full(max(self.n_qubits, 1), -1, dtype=np.int32)
        for q, i in rank.items():
            lut[q] = i
        t = lut[t]
        c = np.where(c >= 0, lut[np.maximum(c, 0)], -1).astype(np.int32)
        identity = np.full(n_physical, -1, dtype=np.int16)
        for q in used:
            if q < n_physical:
                identity[q] = q
        return c, t.astype(np.int32), identity

    def gate_arrays(self):
        """(control, target) int32 arrays; control = -1 for 1q gates."""
        c = np.full(len(self.ops), -1, dtype=np.int32)
        t = np.zeros(len(self.ops), dtype=np.int32)
        for i, op in enumerate(self.ops):
            if len(op.qubits) == 2:
                c[i], t[i] = op.qubits
            else:
                t[i] = op.qubits[0]
        return c, t

    def relabel(self, perm) -> "Circuit":
        """New circuit with logical qubit q renamed perm[q]."""
        out = Circuit(self.n_qubits, name=self.name, n_clbits=self.n_clbits)
        out.ops = [Op(o.name, tuple(int(perm[q]) for q in o.qubits), o.params)
                   for o in self.ops]
        out.measurements = [(int(perm[q]), c) for q, c in self.measurements]
        return out

    def to_qasm(self) -> str:
        lines = ["OPENQASM 2.0;", 'include "qelib1.inc";', f"qreg q[{self.n_qubits}];"]
        if self.n_clbits:
            lines.append(f"creg c[{self.n_clbits}];")
        for op in self.ops:
            p = ""
            if op.params:
                p = "(" + ",".join(repr(float(x)) for x in op.params) + ")"
            lines.append(f"{op.name}{p} " + ",".join(f"q[{q}]" for q in op.qubits) + ";")
        for q, c in self.measurements:
            lines.append(f"measure q[{q}] -> c[{c}];")
        return "\n".join(lines) + "\n"


_TOKEN = re.compile(r"\s*(pi|[0-9]*\.?[0-9]+(?:[eE][-+]?[0-9]+)?|[-+*/()])")


def _eval_expr(s: str) -> float:
    toks = _TOKEN.findall(s)
    if "".join(toks) != re.sub(r"\s+", "", s):
        raise ValueError(f"cannot parse parameter expression: {s!r}")
    pos = 0

    def atom():
        nonlocal pos
        t = toks[pos]
        pos += 1
        if t == "(":
            v = expr()
            pos += 1
            return v
        if t == "-":
            return -atom()
        if t == "+":
            return atom()
        if t == "pi":
            return math.pi
        return float(t)

    def term():
        nonlocal pos
        v = atom()
        while pos < len(toks) and toks[pos] in "*/":
            o = toks[pos]
            pos += 1
            r = atom()
            v = v * r if o == "*" else v / r
        return v

    def expr():
        nonlocal pos
        v = term()
        while pos < len(toks) and toks[pos] in "+-":
            o = toks[pos]
            pos += 1
            r = term()
            v = v + r if o == "+" else v - r
        return v

    return expr()


def _split_params(s: str) -> list:
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch == "," and depth == 0:
            out.append(cur)
            cur = ""
            continue
        depth += ch == "("
        depth -= ch == ")"
        cur += ch
    if cur.strip():
        out.append(cur)
    return out


def parse_qasm(text: str, name: str = "circuit") -> Circuit:
    text = re.sub(r"//[^\n]*", "", text)
    qregs, cregs = {}, {}
    nq = nc = 0
    circ = Circuit(0, name=name)

    def qubit(arg: str) -> int:
        m = re.fullmatch(r"\s*(\w+)\s*\[\s*(\d+)\s*\]\s*", arg)
        if not m or m.group(1) not in qregs:
            raise ValueError(f"bad qubit argument: {arg!r}")
        return qregs[m.group(1)][0] + int(m.group(2))

    for stmt in text.split(";"):
        s = stmt.strip()
        if not s or s.startswith("OPENQASM") or s.startswith("include"):
            continue
        m = re.fullmatch(r"(qreg|creg)\s+(\w+)\s*\[\s*(\d+)\s*\]", s)
        if m:
            size = int(m.group(3))
            if m.group(1) == "qreg":
                qregs[m.group(2)] = (nq, size)
                nq += size
            else:
                cregs[m.group(2)] = (nc, size)
                nc += size
            continue
        if s.startswith("barrier"):
            continue
        if s.startswith("measure"):
            m = re.fullmatch(r"measure\s+(.+?)\s*->\s*(\w+)\s*(?:\[\s*(\d+)\s*\])?", s)
            if not m:
                raise ValueError(f"bad measure: {s!r}")
            if m.group(3) is None:
                reg = re.fullmatch(r"\s*(\w+)\s*", m.group(1)).group(1)
                for k in range(qregs[reg][1]):
                    circ.measurements.append((qregs[reg][0] + k, cregs[m.group(2)][0] + k))
            else:
                circ.measurements.append((qubit(m.group(1)), cregs[m.group(2)][0] + int(m.group(3))))
            continue
        m = re.fullmatch(r"(\w+)\s*(?:\((.*)\))?\s+(.+)", s, flags=re.S)
        if not m:
            raise ValueError(f"cannot parse statement: {s!r}")
        g, params, args = m.group(1), m.group(2), m.group(3)
        qs = [qubit(a) for a in args.split(",")]
        ps = tuple(_eval_expr(p) for p in _split_params(params)) if params else ()
        if g in TWO_QUBIT_GATES:
            if len(qs) != 2:
                raise ValueError(f"bad cx: {s!r}")
            circ.ops.append(Op("cx", (qs[0], qs[1])))
        elif g in ONE_QUBIT_GATES:
            if len(qs) != 1:
                raise ValueError(f"bad 1q gate: {s!r}")
            circ.ops.append(Op(g, (qs[0],), ps))
        elif g == "swap" and len(qs) == 2:
            # only in mapped circuits; the mapper itself wants SWAPs decomposed
            circ.ops.append(Op("swap", (qs[0], qs[1])))
        else:
            # Mapper::createLayers :157-161: gates with more than one control
            raise ValueError(
                "Circuit contains gates with more than one control or an "
                f"unsupported gate ({g}). Please make sure that the input "
                "circuit's gates are decomposed to the appropriate gate set!")
    circ.n_qubits = nq
    circ.n_clbits = nc
    return circ


class NativeQasm:
    """A circuit read by the library's own OpenQASM 2 reader (qmg_qasm_*,
    csrc/qasm_reader.hpp; the host C++ side of large inputs): gate arrays
    without a Python object per gate.  `lib` is the loaded C ABI (_abi.load())."""

    def __init__(self, lib, path: str = None, text: str = None):
        import ctypes as C
        from . import _abi
        self.lib = lib
        h = C.c_void_p()
        if path is not None:
            _abi.check(lib, lib.qmg_qasm_read_file(os.fsencode(path), C.byref(h)))
        else:
            raw = text.encode()
            _abi.check(lib, lib.qmg_qasm_parse(raw, len(raw), C.byref(h)))
        self.h = h
        info = _abi.QmgQasmInfo()
        _abi.check(lib, lib.qmg_qasm_get_info(h, C.byref(info)))
        self.n_qubits, self.n_clbits = int(info.n_qubits), int(info.n_clbits)
        self.n_gates, self.n_measurements = int(info.n_gates), int(info.n_measurements)

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.qmg_qasm_destroy(self.h)
            self.h = None

    def gate_arrays(self):
        """(control, target) as Circuit.gate_arrays()."""
        from . import _abi
        c = np.empty(self.n_gates, dtype=np.int32)
        t = np.empty(self.n_gates, dtype=np.int32)
        _abi.check(self.lib, self.lib.qmg_qasm_gates(self.h, c.ctypes.data, t.ctypes.data, None, None, None))
        return c, t

    def gates(self):
        """(control, target, kind, params[g, 3], n_params) and the kind names."""
        from . import _abi
        g = self.n_gates
        c, t = np.empty(g, np.int32), np.empty(g, np.int32)
        k, npar = np.empty(g, np.uint8), np.empty(g, np.uint8)
        p = np.empty((g, 3), np.float64)
        _abi.check(self.lib, self.lib.qmg_qasm_gates(self.h, c.ctypes.data, t.ctypes.data, k.ctypes.data,
                                                     p.ctypes.data, npar.ctypes.data))
        names = []
        while True:
            nm = self.lib.qmg_qasm_gate_name(len(names))
            if nm is None:
                break
            names.append(nm.decode())
        return c, t, k, p, npar, names

    def measurements(self):
        from . import _abi
        q = np.empty(self.n_measurements, np.uint32)
        c = np.empty(self.n_measurements, np.uint32)
        _abi.check(self.lib, self.lib.qmg_qasm_measurements(self.h, q.ctypes.data, c.ctypes.data))
        return q, c

    def mapper_view(self, n_physical: int):
        """(control, target, identity_locations) as Circuit.mapper_view()."""
        from . import _abi
        c = np.empty(self.n_gates, dtype=np.int32)
        t = np.empty(self.n_gates, dtype=np.int32)
        ident = np.empty(n_physical, dtype=np.int16)
        _abi.check(self.lib, self.lib.qmg_qasm_mapper_view(self.h, n_physical, c.ctypes.data, t.ctypes.data,
                                                           ident.ctypes.data))
        return c, t, ident

    def to_circuit(self, name: str = "circuit") -> Circuit:
        c, t, k, p, npar, names = self.gates()
        circ = Circuit(self.n_qubits, name=name, n_clbits=self.n_clbits)
        for i in range(self.n_gates):
            if c[i] >= 0:
                circ.ops.append(Op(names[k[i]], (int(c[i]), int(t[i]))))
            else:
                circ.ops.append(Op(names[k[i]], (int(t[i]),), tuple(float(x) for x in p[i, :npar[i]])))
        q, cl = self.measurements()
        circ.measurements = [(int(a), int(b)) for a, b in zip(q, cl)]
        return circ


def load_qasm_native(lib, path: str) -> NativeQasm:
    """The file through the library's reader (fails like parse_qasm does)."""
    return NativeQasm(lib, path=path)


def load_qasm(path: str) -> Circuit:
    import os
    with open(path) as f:
        return parse_qasm(f.read(), name=os.path.splitext(os.path.basename(path))[0])


def random_circuit(n_qubits: int, n_cx: int, seed: int, p_1q: float = 0.5) -> Circuit:
    """Seeded synthetic circuit of BASELINE.md (configs 4/5): n_cx CX on
    uniformly random distinct pairs, one random 1q gate (h/t/x) in between
    with probability p_1q."""
    rng = np.random.default_rng(seed)
    c = Circuit(n_qubits, name=f"random_{n_qubits}q_{n_cx}cx_s{seed}")
    a = rng.integers(0, n_qubits, size=n_cx)
    b = rng.integers(0, n_qubits - 1, size=n_cx)
    b = np.where(b >= a, b + 1, b)
    one = rng.random(n_cx) < p_1q
    oneq = rng.integers(0, n_qubits, size=n_cx)
    kind = rng.integers(0, 3, size=n_cx)
    for i in range(n_cx):
        c.cx(int(a[i]), int(b[i]))
        if one[i]:
            c.gate(("h", "t", "x")[kind[i]], int(oneq[i]))
    return c
