"""Architecture: coupling map + calibration data + the device-resident cost
tables.  Mirrors class Architecture (include/sc/Architecture.hpp:42,
src/sc/Architecture.cpp): the parsers and the -log2 fidelity costs stay on the
host exactly as in the reference (Architecture.cpp:52-83, :108-174, :232-289);
the distance / fidelity-distance tables are built by the CUDA kernels behind
qmg_arch_create / qmg_arch_set_fidelity_costs and copied back for the getters.
"""
from __future__ import annotations

import ctypes as C
import math
import os
import re

import numpy as np

from . import _abi
from ._abi import QmapError

DBL_MAX = float(np.finfo(np.float64).max)

# include/sc/configuration/AvailableArchitecture.hpp:88-120
AVAILABLE_ARCHITECTURES = {
    "IBM_QX4": "5\n1 0\n2 0\n2 1\n3 2\n3 4\n2 4",
    "IBM_QX5": "16\n1 0\n15 0\n1 2\n2 3\n15 2\n3 4\n3 14\n5 4\n13 4\n6 5\n12 5\n6 7\n6 11\n8 7\n"
               "7 10\n9 8\n9 10\n11 10\n12 11\n12 13\n13 14\n15 14",
    "IBMQ_Yorktown": "5\n0 1\n1 0\n0 2\n2 0\n1 2\n2 1\n2 3\n3 2\n3 4\n4 3\n2 4\n4 2",
    "IBMQ_London": "5\n0 1\n1 0\n1 2\n2 1\n1 3\n3 1\n3 4\n4 3",
    "IBMQ_Bogota": "5\n0 1\n1 0\n1 2\n2 1\n2 3\n3 2\n3 4\n4 3",
    "IBMQ_Casablanca": "7\n0 1\n1 0\n1 2\n2 1\n1 3\n3 1\n3 5\n5 3\n5 4\n4 5\n5 6\n6 5",
    "Rigetti_Agave": "8\n1 0\n0 1\n0 7\n7 0\n7 6\n6 7\n6 5\n5 6\n5 4\n4 5\n4 3\n3 4\n3 2\n2 3\n2 1\n1 2",
    "Rigetti_Aspen": "16\n0 1\n1 0\n1 14\n14 1\n14 15\n15 14\n15 0\n0 15\n0 7\n7 0\n7 6\n6 7\n6 5\n"
                     "5 6\n5 4\n4 5\n4 3\n3 4\n3 2\n2 3\n2 1\n1 2\n14 13\n13 14\n13 12\n12 13\n"
                     "12 11\n11 12\n11 10\n10 11\n10 9\n9 10\n9 8\n8 9\n8 15\n15 8",
}


def _tokyo() -> str:
    und = [(0, 1), (1, 2), (2, 3), (3, 4), (5, 6), (6, 7), (7, 8), (8, 9), (10, 11), (11, 12),
           (12, 13), (13, 14), (15, 16), (16, 17), (17, 18), (18, 19), (0, 5), (5, 10), (10, 15),
           (1, 6), (6, 11), (11, 16), (2, 7), (7, 12), (12, 17), (3, 8), (8, 13), (13, 18),
           (4, 9), (9, 14), (14, 19), (5, 11), (11, 17), (1, 7), (7, 13), (13, 9), (3, 9),
           (2, 6), (6, 10), (4, 8), (8, 12), (12, 16), (14, 18)]
    return "20\n" + "\n".join(f"{a} {b}\n{b} {a}" for a, b in und)


AVAILABLE_ARCHITECTURES["IBMQ_Tokyo"] = _tokyo()


# ---------------------------------------------------------------------------
# synthetic architectures of BASELINE.json's configs 3-5
# ---------------------------------------------------------------------------
def _bidir(und):
    out = set()
    for a, b in und:
        out.add((a, b))
        out.add((b, a))
    return sorted(out)


def heavy_hex_27():
    """27-qubit heavy-hex (IBM Falcon coupling list), bidirectional."""
    und = [(0, 1), (1, 2), (1, 4), (2, 3), (3, 5), (4, 7), (5, 8), (6, 7), (7, 10), (8, 9),
           (8, 11), (10, 12), (11, 14), (12, 13), (12, 15), (13, 14), (14, 16), (15, 18),
           (16, 19), (17, 18), (18, 21), (19, 20), (19, 22), (21, 23), (22, 25), (23, 24),
           (24, 25), (25, 26)]
    return 27, _bidir(und)


def heavy_hex_127():
    """127-qubit heavy-hex (IBM Eagle layout): 7 rows of 14/15 qubits joined by
    4 bridge qubits between consecutive rows; 144 undirected edges."""
    rows = [(0, 14, 0)]  # (first qubit, length, column of first qubit)
    q = 14
    und = []
    bridges = []
    for r in range(1, 7):
        bridges.append(list(range(q, q + 4)))
        q += 4
        length = 14 if r == 6 else 15
        rows.append((q, length, 1 if r == 6 else 0))
        q += length
    assert q == 127
    for first, length, _ in rows:
        und += [(first + i, first + i + 1) for i in range(length - 1)]
    for r, br in enumerate(bridges):
        cols = (0, 4, 8, 12) if r % 2 == 0 else (2, 6, 10, 14)
        f0, _, c0 = rows[r]
        f1, _, c1 = rows[r + 1]
        for bq, col in zip(br, cols):
            und.append((f0 + col - c0, bq))
            und.append((bq, f1 + col - c1))
    assert len(und) == 144
    return 127, _bidir(und)


def grid(rows: int, cols: int):
    und = []
    for r in range(rows):
        for c in range(cols):
            v = r * cols + c
            if c + 1 < cols:
                und.append((v, v + 1))
            if r + 1 < rows:
                und.append((v, v + cols))
    return rows * cols, _bidir(und)


def line(n: int, bidirectional=True):
    und = [(i, i + 1) for i in range(n - 1)]
    return n, (_bidir(und) if bidirectional else sorted(und))


def synthetic_calibration_csv(n: int, coupling_map, seed: int) -> str:
    """Seeded calibration data in the reference's CSV format
    (extern/calibration/ibmq_london.csv): 1q error ~ logU[1e-4, 1e-3], CX error
    ~ logU[5e-3, 3e-2], one rate per undirected edge."""
    rng = np.random.default_rng(seed)
    q1 = np.exp(rng.uniform(math.log(1e-4), math.log(1e-3), size=n))
    und = sorted({(min(a, b), max(a, b)) for a, b in coupling_map})
    cx = dict(zip(und, np.exp(rng.uniform(math.log(5e-3), math.log(3e-2), size=len(und)))))
    lines = ["Qubit,T1 (µs),T2 (µs),Frequency (GHz),Readout error,"
             "Single-qubit U2 error rate,CNOT error rate,Date"]
    cm = set(coupling_map)
    for q in range(n):
        outs = sorted(b for a, b in cm if a == q)
        cxs = ", ".join(f"cx{q}_{b}: {float(cx[(min(q, b), max(q, b))])!r}" for b in outs)
        lines.append(f"Q{q},50.0,60.0,5.0,2.0e-2,{float(q1[q])!r},\"{cxs}\",synthetic seed {seed}")
    return "\n".join(lines) + "\n"


# ---------------------------------------------------------------------------
class Properties:
    """Architecture::Properties restricted to what the fidelity tables read."""
    SINGLE_QUBIT_GATES = ("id", "u1", "u2", "u3", "rz", "sx", "x")

    def __init__(self):
        self.name = ""
        self.n_qubits = 0
        self.single_qubit_error = {}  # qubit -> {op: rate}
        self.two_qubit_error = {}  # (a, b) -> rate
        self.readout_error = {}
        self.t1 = {}
        self.t2 = {}
        self.frequency = {}
        self.calibration_date = {}

    def set_single_qubit_error_rate(self, q, op, rate):
        self.single_qubit_error.setdefault(int(q), {})[op] = float(rate)

    def set_two_qubit_error_rate(self, a, b, rate):
        self.two_qubit_error[(int(a), int(b))] = float(rate)

    def average_single_qubit_error_rate(self, q) -> float:
        # Architecture.hpp:101-109 (same summation order: one term per op)
        ops = self.single_qubit_error[q]
        avg = 0.0
        for _, e in sorted(ops.items()):
            avg += e
        return avg / float(len(ops))

    def empty(self) -> bool:
        return not self.single_qubit_error and not self.two_qubit_error


def _parse_csv_line(line: str) -> list:
    out, cur, quoted = [], "", False
    for ch in line:
        if ch == '"':
            quoted = not quoted
        elif ch == "," and not quoted:
            out.append(cur)
            cur = ""
        else:
            cur += ch
    out.append(cur)
    return out


_CX_RE = re.compile(r"((\d+).?(\d+):\W*?(-?(?:0|[1-9]\d*)(?:\.\d+)?(?:[eE][+-]?\d+)?))")


class Architecture:
    def __init__(self, n_qubits: int | None = None, coupling_map=None, properties=None,
                 device: int = 0):
        self.name = ""
        self.n_qubits = 0
        self.coupling_map = []  # sorted list of (first, second): std::set order
        self.properties = Properties()
        self.device = device
        self._h = None
        self._lib = None
        self._dist = self._dist_rev = self._fid = None
        self._fid_costs = None
        self.fidelity_available = False
        if n_qubits is not None:
            self.load_coupling_map(n_qubits, coupling_map or [])
            if properties is not None:
                self.load_properties(properties)

    # ---- coupling map --------------------------------------------------
    def load_coupling_map(self, spec, coupling_map=None):
        """spec: qubit count (with coupling_map), a built-in name, a path or
        .arch text."""
        self.properties = Properties()
        self.fidelity_available = False
        if isinstance(spec, (int, np.integer)):
            self.n_qubits = int(spec)
            self.coupling_map = sorted({(int(a), int(b)) for a, b in coupling_map})
            self.name = f"generic_{self.n_qubits}"
        else:
            if spec in AVAILABLE_ARCHITECTURES:
                text, self.name = AVAILABLE_ARCHITECTURES[spec], spec
            elif os.path.exists(spec):
                with open(spec) as f:
                    text = f.read()
                self.name = os.path.splitext(os.path.basename(spec))[0]
            elif "\n" in spec:
                text = spec
            else:
                raise QmapError("Error opening coupling map file.")
            self._parse_arch(text)
        for a, b in self.coupling_map:
            if a >= self.n_qubits or b >= self.n_qubits:
                raise QmapError("coupling map entry out of range")
        self._invalidate()
        return self

    def _parse_arch(self, text: str):
        # Architecture.cpp:52-83
        lines = text.split("\n")
        while lines and lines[-1] == "":
            lines.pop()
        if not lines:
            raise QmapError("Error reading coupling map file.")
        m = re.search(r"([0-9]+)", lines[0])
        if not m:
            raise QmapError("No qubit count found in coupling map file: " + lines[0])
        self.n_qubits = int(m.group(1))
        cm = set()
        for ln in lines[1:]:
            m = re.search(r"([0-9]+) ([0-9]+)", ln)
            if not m:
                raise QmapError("Could not identify edge in coupling map file: " + ln)
            cm.add((int(m.group(1)), int(m.group(2))))
        self.coupling_map = sorted(cm)

    # ---- calibration ---------------------------------------------------
    def load_properties(self, spec):
        """spec: Properties, a path or CSV text (Architecture.cpp:94-189)."""
        if isinstance(spec, Properties):
            props = spec
            # Architecture.cpp:176-186: the coupling map follows from the
            # properties only when no architecture is loaded (name and qubit
            # count, isArchitectureAvailable)
            if not (self.name and self.n_qubits):
                self.coupling_map = sorted(props.two_qubit_error)
                self.n_qubits = props.n_qubits
                self.name = f"generic_{self.n_qubits}"
                self._invalidate()
        else:
            if os.path.exists(spec):
                with open(spec) as f:
                    text = f.read()
                pname = os.path.splitext(os.path.basename(spec))[0]
            elif "\n" in spec:
                text, pname = spec, "properties"
            else:
                raise QmapError("Error opening properties file.")
            props = self._parse_csv(text)
            props.name = pname
            if not self.name:
                self.name = pname
        self.properties = props
        self._create_fidelity_costs()
        return self

    def _parse_csv(self, text: str) -> Properties:
        props = Properties()
        have_arch = bool(self.coupling_map)
        cm = set(self.coupling_map)
        avg, cnt = 0.0, 0
        q = 0
        lines = text.split("\n")[1:]
        for ln in lines:
            if ln.strip() == "":
                continue
            data = _parse_csv_line(ln)
            props.t1[q] = float(data[1])
            props.t2[q] = float(data[2])
            props.frequency[q] = float(data[3])
            props.readout_error[q] = float(data[4])
            for op in Properties.SINGLE_QUBIT_GATES:
                props.set_single_qubit_error_rate(q, op, float(data[5]))
            if len(data) >= 7:
                for m in _CX_RE.finditer(data[6]):
                    a, b, r = int(m.group(2)), int(m.group(3)), float(m.group(4))
                    if not have_arch:
                        cm.add((a, b))
                    cnt += 1
                    avg = avg + (r - avg) / cnt
                    props.set_two_qubit_error_rate(a, b, r)
            if len(data) == 8:
                props.calibration_date[q] = data[7]
            q += 1
        if have_arch:
            for e in self.coupling_map:
                if e not in props.two_qubit_error:
                    props.set_two_qubit_error_rate(e[0], e[1], avg)
        props.n_qubits = q
        if not have_arch:
            self.n_qubits = q
            self.coupling_map = sorted(cm)
            self._invalidate()
        return props

    def _create_fidelity_costs(self):
        """Host part of Architecture::createFidelityTable (:232-289)."""
        n = self.n_qubits
        p = self.properties
        self.fidelity_available = True
        c1 = np.zeros(n)
        c2 = np.full((n, n), DBL_MAX)
        cs = np.full((n, n), DBL_MAX)
        for q in sorted(p.single_qubit_error):
            fid = 1.0 - p.average_single_qubit_error_rate(q)
            c1[q] = -math.log2(fid)
        cm = set(self.coupling_map)
        for a, b in self.coupling_map:
            if (a, b) not in p.two_qubit_error:
                self.fidelity_available = False
                self._fid_costs = None
                self._fid = None
                return
            c2[a, b] = -math.log2(1.0 - p.two_qubit_error[(a, b)])
            if (b, a) not in cm:
                c2[b, a] = c2[a, b] + 2 * c1[a] + 2 * c1[b]
                cs[a, b] = 3 * c2[a, b] + 2 * c1[a] + 2 * c1[b]
                cs[b, a] = cs[a, b]
            else:
                cs[a, b] = 3 * c2[a, b]
        self._fid_costs = (c1, c2, cs)
        self._fid = None
        if self._h is not None:
            self._upload_fidelity()

    # ---- device handle -------------------------------------------------
    def _invalidate(self):
        if self._h is not None:
            self._lib.qmg_arch_destroy(self._h)
        self._h = None
        self._dist = self._dist_rev = self._fid = None

    def __del__(self):
        try:
            self._invalidate()
        except Exception:
            pass

    @property
    def edges(self) -> np.ndarray:
        return np.asarray(self.coupling_map, dtype=np.uint16).reshape(-1, 2)

    @property
    def handle(self):
        """qmg_arch*: created on first use (tables are built on the GPU)."""
        if self._h is None:
            self._lib = L = _abi.load()
            e = np.ascontiguousarray(self.edges)
            h = C.c_void_p()
            _abi.check(L, L.qmg_arch_create(self.n_qubits, e.ctypes.data, len(e),
                                            self.device, C.byref(h)))
            self._h = h
            if self._fid_costs is not None:
                self._upload_fidelity()
        return self._h

    def _upload_fidelity(self):
        c1, c2, cs = (np.ascontiguousarray(x) for x in self._fid_costs)
        _abi.check(self._lib, self._lib.qmg_arch_set_fidelity_costs(
            self._h, c1.ctypes.data, c2.ctypes.data, cs.ctypes.data))

    def info(self) -> _abi.QmgArchInfo:
        i = _abi.QmgArchInfo()
        _abi.check(self._lib or _abi.load(), _abi.load().qmg_arch_get_info(self.handle, C.byref(i)))
        return i

    # ---- getters (Architecture.hpp) ------------------------------------
    def bidirectional(self) -> bool:
        cm = set(self.coupling_map)
        return all((b, a) in cm for a, b in cm)

    def unidirectional(self) -> bool:
        cm = set(self.coupling_map)
        return all((b, a) not in cm for a, b in cm)

    def is_edge_connected(self, a, b, consider_direction=True) -> bool:
        cm = self._cm_set()
        return (a, b) in cm or (not consider_direction and (b, a) in cm)

    def _cm_set(self):
        s = getattr(self, "_cm_cache", None)
        if s is None or self._cm_cache_src is not self.coupling_map:
            self._cm_cache = s = set(self.coupling_map)
            self._cm_cache_src = self.coupling_map
        return s

    def distance_table(self, include_reversal_cost=True) -> np.ndarray:
        if self._dist is None:
            h = self.handle
            n = self.n_qubits
            self._dist = np.empty((n, n))
            self._dist_rev = np.empty((n, n))
            L = self._lib
            _abi.check(L, L.qmg_arch_get_distance_table(h, 0, self._dist.ctypes.data))
            _abi.check(L, L.qmg_arch_get_distance_table(h, 1, self._dist_rev.ctypes.data))
        return self._dist_rev if include_reversal_cost else self._dist

    def distance(self, a, b, include_reversal_cost=True) -> float:
        return float(self.distance_table(include_reversal_cost)[a, b])

    def _need_fidelity(self):
        if not self.fidelity_available:
            raise QmapError("No fidelity data available.")

    def single_qubit_fidelity_costs(self):
        self._need_fidelity()
        return self._fid_costs[0]

    def two_qubit_fidelity_costs(self):
        self._need_fidelity()
        return self._fid_costs[1]

    def swap_fidelity_costs(self):
        self._need_fidelity()
        return self._fid_costs[2]

    def fidelity_distance_tables(self) -> np.ndarray:
        self._need_fidelity()
        if self._fid is None:
            h = self.handle
            k = int(self.info().n_fidelity_tables)
            n = self.n_qubits
            self._fid = np.empty((k, n, n))
            _abi.check(self._lib, self._lib.qmg_arch_get_fidelity_distance_tables(
                h, self._fid.ctypes.data))
        return self._fid

    def fidelity_distance(self, a, b, skip_edges=0) -> float:
        t = self.fidelity_distance_tables()
        if skip_edges >= len(t):
            return 0.0
        return float(t[skip_edges, a, b])

    # ---- utilities next to the tables (host only, used by the
    # reference's exact mapper, nothing on the heuristic path calls them)
    def minimum_number_of_swaps(self, permutation, limit: int = -1) -> int:
        """Architecture::minimumNumberOfSwaps(permutation, limit)
        (src/sc/Architecture.cpp:296-385): the fewest SWAPs on coupling edges
        among the qubits of `permutation` that take the identity to the goal
        "sorted(permutation)[k] holds permutation[k]"; with a limit, limit + 1
        as soon as no solution below it exists.  Breadth-first over layouts
        with a visited set (the reference's uniform-cost search without one
        returns the same count)."""
        perm = [int(q) for q in permutation]
        qubits = sorted(set(perm))
        goal = {q: perm[k] for k, q in enumerate(qubits)}
        if all(q == g for q, g in goal.items()):
            return 0
        qs = set(qubits)
        cm = self._cm_set()
        swaps = set()
        for a, b in self.coupling_map:
            if a in qs and b in qs and (not self.bidirectional() or ((a, b) not in swaps and (b, a) not in swaps)):
                swaps.add((a, b))
        swaps = sorted(swaps)
        start = tuple(range(self.n_qubits))
        frontier = [start]
        seen = {start}
        depth = 0
        while frontier:
            if limit != -1 and depth >= limit:
                return limit + 1
            nxt = []
            for cur in frontier:
                for a, b in swaps:
                    lst = list(cur)
                    lst[a], lst[b] = lst[b], lst[a]
                    if all(lst[q] == g for q, g in goal.items()):
                        return depth + 1
                    t = tuple(lst)
                    if t not in seen:
                        seen.add(t)
                        nxt.append(t)
            frontier = nxt
            depth += 1
        return 0

    def get_coupling_limit(self, qubit_choice=None) -> int:
        """Architecture::getCouplingLimit() / getCouplingLimit(qubitChoice)
        (src/sc/Architecture.cpp:480-569): the longest of the shortest paths in
        the undirected coupling graph (induced on qubit_choice): the
        reference's exhaustive walk over simple paths keeps the smallest step
        count per qubit, i.e. its breadth-first distance."""
        choice = None if qubit_choice is None else set(int(q) for q in qubit_choice)
        adj = [set() for _ in range(self.n_qubits)]
        for a, b in self.coupling_map:
            if choice is None or (a in choice and b in choice):
                adj[a].add(b)
                adj[b].add(a)
        best = 0
        for s0 in range(self.n_qubits):
            if choice is not None and not adj[s0]:
                continue
            dist = {s0: 0}
            frontier = [s0]
            while frontier:
                nxt = []
                for u in frontier:
                    for v in adj[u]:
                        if v not in dist:
                            dist[v] = dist[u] + 1
                            nxt.append(v)
                frontier = nxt
            best = max(best, max(dist.values()))
        return best

    def to_arch_text(self) -> str:
        return f"{self.n_qubits}\n" + "\n".join(f"{a} {b}" for a, b in self.coupling_map) + "\n"
