"""ctypes bindings for the test-infrastructure libraries under oracle/.

TEST INFRASTRUCTURE ONLY.  `Ref` wraps oracle/_ref/libqmapref.so (the
reference's own sources, see ref_driver.cpp); `Oracle` (oracle/oracleapi.py)
wraps the CPU restatement.  Only tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py import this module.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(HERE, "_ref", "libqmapref.so")


class QmgConfig(C.Structure):
    """Mirror of qmg_config (include/qmap_b200.h)."""

    _fields_ = [
        ("heuristic", C.c_int32),
        ("lookahead_heuristic", C.c_int32),
        ("nr_lookaheads", C.c_uint32),
        ("_pad0", C.c_uint32),
        ("first_lookahead_factor", C.c_double),
        ("lookahead_factor", C.c_double),
        ("early_termination", C.c_int32),
        ("automatic_layer_splits", C.c_int32),
        ("early_termination_limit", C.c_uint64),
        ("automatic_layer_splits_node_limit", C.c_uint64),
        ("layering", C.c_int32),
        ("initial_layout", C.c_int32),
        ("iterative_bidirectional_routing_passes", C.c_uint32),
        ("swap_on_first_layer", C.c_int32),
        ("add_barriers_between_layers", C.c_int32),
        ("add_measurements_to_mapped_circuit", C.c_int32),
        ("pre_mapping_optimizations", C.c_int32),
        ("post_mapping_optimizations", C.c_int32),
        ("debug", C.c_int32),
        ("verbose", C.c_int32),
    ]


HEURISTICS = {
    "gate_count_max_distance": 0,
    "gate_count_sum_distance": 1,
    "gate_count_sum_distance_minus_shared_swaps": 2,
    "gate_count_max_distance_or_sum_distance_minus_shared_swaps": 3,
    "fidelity_best_location": 4,
}
LOOKAHEADS = {"none": 0, "gate_count_max_distance": 1, "gate_count_sum_distance": 2}
LAYERINGS = {
    "individual_gates": 0,
    "disjoint_qubits": 1,
    "odd_gates": 2,
    "qubit_triangle": 3,
    "disjoint_2q_blocks": 4,
}
LAYOUTS = {"identity": 0, "static": 1, "dynamic": 2}
EARLY_TERMINATIONS = {
    "none": 0,
    "expanded_nodes": 1,
    "expanded_nodes_after_first_solution": 2,
    "expanded_nodes_after_current_optimal_solution": 3,
    "solution_nodes": 4,
    "solution_nodes_after_current_optimal_solution": 5,
}


def make_config(
    heuristic="gate_count_max_distance",
    lookahead_heuristic="gate_count_max_distance",
    nr_lookaheads=15,
    first_lookahead_factor=0.75,
    lookahead_factor=0.5,
    early_termination="none",
    early_termination_limit=0,
    automatic_layer_splits=True,
    automatic_layer_splits_node_limit=5000,
    layering="individual_gates",
    initial_layout="dynamic",
    iterative_bidirectional_routing_passes=0,
    swap_on_first_layer=False,
    add_barriers_between_layers=False,
    add_measurements_to_mapped_circuit=True,
    pre_mapping_optimizations=True,
    post_mapping_optimizations=True,
    debug=True,
    verbose=False,
) -> QmgConfig:
    """Configuration with the reference's defaults
    (include/sc/configuration/Configuration.hpp:30-119)."""
    c = QmgConfig()
    c.heuristic = HEURISTICS[heuristic] if isinstance(heuristic, str) else int(heuristic)
    c.lookahead_heuristic = (
        LOOKAHEADS[lookahead_heuristic]
        if isinstance(lookahead_heuristic, str)
        else int(lookahead_heuristic)
    )
    c.nr_lookaheads = nr_lookaheads
    c.first_lookahead_factor = first_lookahead_factor
    c.lookahead_factor = lookahead_factor
    c.early_termination = (
        EARLY_TERMINATIONS[early_termination]
        if isinstance(early_termination, str)
        else int(early_termination)
    )
    c.early_termination_limit = early_termination_limit
    c.automatic_layer_splits = int(automatic_layer_splits)
    c.automatic_layer_splits_node_limit = automatic_layer_splits_node_limit
    c.layering = LAYERINGS[layering] if isinstance(layering, str) else int(layering)
    c.initial_layout = LAYOUTS[initial_layout] if isinstance(initial_layout, str) else int(initial_layout)
    c.iterative_bidirectional_routing_passes = iterative_bidirectional_routing_passes
    c.swap_on_first_layer = int(swap_on_first_layer)
    c.add_barriers_between_layers = int(add_barriers_between_layers)
    c.add_measurements_to_mapped_circuit = int(add_measurements_to_mapped_circuit)
    c.pre_mapping_optimizations = int(pre_mapping_optimizations)
    c.post_mapping_optimizations = int(post_mapping_optimizations)
    c.debug = int(debug)
    c.verbose = int(verbose)
    return c


class QmgLayerResult(C.Structure):
    """Mirror of qmg_layer_result (include/qmap_b200.h)."""

    _fields_ = [
        ("cost_fixed", C.c_double),
        ("cost_fixed_reversals", C.c_double),
        ("cost_heur", C.c_double),
        ("lookahead_penalty", C.c_double),
        ("expanded_nodes", C.c_uint64),
        ("generated_nodes", C.c_uint64),
        ("expanded_nodes_after_first_solution", C.c_uint64),
        ("expanded_nodes_after_optimal_solution", C.c_uint64),
        ("solution_nodes", C.c_uint64),
        ("solution_nodes_after_optimal_solution", C.c_uint64),
        ("layout_hash", C.c_uint64),
        ("solution_depth", C.c_uint32),
        ("result_node_id", C.c_uint32),
        ("swap_offset", C.c_uint32),
        ("n_swaps", C.c_uint32),
        ("placement_offset", C.c_uint32),
        ("n_placements", C.c_uint32),
        ("early_termination", C.c_int32),
        ("status", C.c_int32),
    ]


LAYER_RESULT_DTYPE = np.dtype(
    [
        ("cost_fixed", "f8"),
        ("cost_fixed_reversals", "f8"),
        ("cost_heur", "f8"),
        ("lookahead_penalty", "f8"),
        ("expanded_nodes", "u8"),
        ("generated_nodes", "u8"),
        ("expanded_nodes_after_first_solution", "u8"),
        ("expanded_nodes_after_optimal_solution", "u8"),
        ("solution_nodes", "u8"),
        ("solution_nodes_after_optimal_solution", "u8"),
        ("layout_hash", "u8"),
        ("solution_depth", "u4"),
        ("result_node_id", "u4"),
        ("swap_offset", "u4"),
        ("n_swaps", "u4"),
        ("placement_offset", "u4"),
        ("n_placements", "u4"),
        ("early_termination", "i4"),
        ("status", "i4"),
    ]
)
assert LAYER_RESULT_DTYPE.itemsize == C.sizeof(QmgLayerResult)


class QmgArchInfo(C.Structure):
    _fields_ = [
        ("n", C.c_uint16),
        ("_pad", C.c_uint16),
        ("n_edges", C.c_uint32),
        ("bidirectional", C.c_int32),
        ("unidirectional", C.c_int32),
        ("fidelity_available", C.c_int32),
        ("n_fidelity_tables", C.c_uint32),
        ("device", C.c_int32),
        ("table_build_ms", C.c_float),
    ]


class _RefSummary(C.Structure):
    _fields_ = [
        ("seconds", C.c_double),
        ("results_time", C.c_double),
        ("seconds_per_node", C.c_double),
        ("expanded_nodes", C.c_uint64),
        ("generated_nodes", C.c_uint64),
        ("swaps", C.c_uint64),
        ("direction_reverse", C.c_uint64),
        ("output_gates", C.c_uint64),
        ("output_cnots", C.c_uint64),
        ("output_single_qubit_gates", C.c_uint64),
        ("layers", C.c_uint64),
        ("n_calls", C.c_uint64),
        ("n_swap_pairs", C.c_uint64),
        ("n_mapped_ops", C.c_uint64),
        ("total_log_fidelity", C.c_double),
    ]


def _edges_array(edges) -> np.ndarray:
    e = np.asarray(sorted({(int(a), int(b)) for a, b in edges}), dtype=np.uint16)
    return np.ascontiguousarray(e.reshape(-1, 2))


def _p(a: np.ndarray, t):
    return a.ctypes.data_as(C.POINTER(t))


@dataclass
class RefMapResult:
    """Everything the reference's HeuristicMapper::map produced for one run."""

    seconds: float
    results_time: float
    seconds_per_node: float
    expanded_nodes: int
    generated_nodes: int
    swaps: int
    direction_reverse: int
    output_gates: int
    output_cnots: int
    output_single_qubit_gates: int
    n_layers: int
    total_log_fidelity: float
    calls: np.ndarray  # LAYER_RESULT_DTYPE, one per aStarMap call
    call_layer: np.ndarray
    call_flags: np.ndarray  # bit0 reverse, bit1 split
    qubits_before: np.ndarray  # (n_calls, n)
    qubits_after: np.ndarray
    locations_after: np.ndarray
    swap_pairs: np.ndarray  # (total, 2)
    mapped_ops: np.ndarray  # (n_ops, 3): (OpType, q0, q1)
    initial_layout: np.ndarray  # physical -> logical (-1 none)
    output_permutation: np.ndarray
    final_qubits: np.ndarray
    json: str


class Ref:
    """The compiled reference (oracle/_ref/libqmapref.so)."""

    def __init__(self, path: str = REF_LIB):
        if not os.path.exists(path):
            raise FileNotFoundError(
                f"{path} missing: run `make -C oracle ref` where /root/reference exists"
            )
        self.lib = L = C.CDLL(path)
        L.qref_last_error.restype = C.c_char_p
        L.qref_arch_create.restype = C.c_void_p
        L.qref_arch_create.argtypes = [C.c_uint16, C.c_void_p, C.c_size_t]
        L.qref_arch_load.restype = C.c_void_p
        L.qref_arch_load.argtypes = [C.c_char_p, C.c_char_p]
        L.qref_arch_set_error_rates.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.qref_arch_destroy.argtypes = [C.c_void_p]
        L.qref_arch_info.argtypes = [C.c_void_p, C.POINTER(QmgArchInfo)]
        L.qref_arch_edges.argtypes = [C.c_void_p, C.c_void_p]
        L.qref_arch_distance_table.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.qref_arch_fidelity.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.qref_build_table.argtypes = [C.c_uint16, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
        L.qref_build_single_edge_skip_table.argtypes = [
            C.c_uint16, C.c_void_p, C.c_void_p, C.c_size_t, C.c_double, C.c_void_p]
        L.qref_build_edge_skip_table.argtypes = [
            C.c_uint16, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_size_t,
            C.POINTER(C.c_size_t)]
        L.qref_map_qasm.restype = C.c_void_p
        L.qref_map_qasm.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(QmgConfig), C.c_char_p]
        L.qref_map_gates.restype = C.c_void_p
        L.qref_map_gates.argtypes = [
            C.c_void_p, C.c_uint32, C.c_size_t, C.c_void_p, C.c_void_p,
            C.POINTER(QmgConfig), C.c_char_p]
        L.qref_result_destroy.argtypes = [C.c_void_p]
        L.qref_result_summary.argtypes = [C.c_void_p, C.POINTER(_RefSummary)]
        L.qref_result_calls.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.qref_result_layouts.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.qref_result_circuit.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.qref_result_json.restype = C.c_char_p
        L.qref_result_json.argtypes = [C.c_void_p]

    def _err(self) -> str:
        return self.lib.qref_last_error().decode()

    # ---- Dijkstra:: ----
    def build_table(self, n, edges, weights):
        e = _edges_array(edges)
        w = np.ascontiguousarray(weights, dtype=np.float64)
        out = np.empty((n, n), dtype=np.float64)
        if self.lib.qref_build_table(n, e.ctypes.data, len(e), w.ctypes.data, out.ctypes.data):
            raise RuntimeError(self._err())
        return out

    def build_single_edge_skip_table(self, dist, edges, reversal_cost):
        d = np.ascontiguousarray(dist, dtype=np.float64)
        n = d.shape[0]
        e = _edges_array(edges)
        out = np.empty((n, n), dtype=np.float64)
        if self.lib.qref_build_single_edge_skip_table(
                n, d.ctypes.data, e.ctypes.data, len(e), float(reversal_cost), out.ctypes.data):
            raise RuntimeError(self._err())
        return out

    def build_edge_skip_table(self, n, edges, weights, max_tables=None):
        e = _edges_array(edges)
        w = np.ascontiguousarray(weights, dtype=np.float64)
        max_tables = max_tables or (n + 1)
        out = np.zeros((max_tables, n, n), dtype=np.float64)
        k = C.c_size_t(0)
        if self.lib.qref_build_edge_skip_table(
                n, e.ctypes.data, len(e), w.ctypes.data, out.ctypes.data, max_tables, C.byref(k)):
            raise RuntimeError(self._err())
        return out[: k.value].copy()

    # ---- Architecture ----
    def arch_create(self, n, edges) -> "RefArch":
        e = _edges_array(edges)
        h = self.lib.qref_arch_create(n, e.ctypes.data, len(e))
        if not h:
            raise RuntimeError(self._err())
        return RefArch(self, h)

    def arch_load(self, arch_file: str, calibration_file: str | None = None) -> "RefArch":
        h = self.lib.qref_arch_load(
            (arch_file or "").encode(), (calibration_file or "").encode())
        if not h:
            raise RuntimeError(self._err())
        return RefArch(self, h)

    # ---- HeuristicMapper::map ----
    def _collect(self, arch: "RefArch", h) -> RefMapResult:
        if not h:
            raise RuntimeError(self._err())
        try:
            s = _RefSummary()
            self.lib.qref_result_summary(h, C.byref(s))
            n = arch.n
            nc = int(s.n_calls)
            calls = np.zeros(nc, dtype=LAYER_RESULT_DTYPE)
            layer = np.zeros(nc, dtype=np.uint32)
            flags = np.zeros(nc, dtype=np.int32)
            self.lib.qref_result_calls(h, calls.ctypes.data, layer.ctypes.data, flags.ctypes.data)
            qb = np.zeros((nc, n), dtype=np.int16)
            qa = np.zeros((nc, n), dtype=np.int16)
            la = np.zeros((nc, n), dtype=np.int16)
            sw = np.zeros((int(s.n_swap_pairs), 2), dtype=np.uint16)
            self.lib.qref_result_layouts(h, qb.ctypes.data, qa.ctypes.data, la.ctypes.data, sw.ctypes.data)
            ops = np.zeros((int(s.n_mapped_ops), 3), dtype=np.int32)
            il = np.zeros(n, dtype=np.int32)
            op = np.zeros(n, dtype=np.int32)
            fq = np.zeros(n, dtype=np.int16)
            self.lib.qref_result_circuit(h, ops.ctypes.data, il.ctypes.data, op.ctypes.data, fq.ctypes.data)
            js = self.lib.qref_result_json(h).decode()
            return RefMapResult(
                s.seconds, s.results_time, s.seconds_per_node, int(s.expanded_nodes),
                int(s.generated_nodes), int(s.swaps), int(s.direction_reverse),
                int(s.output_gates), int(s.output_cnots), int(s.output_single_qubit_gates),
                int(s.layers), s.total_log_fidelity, calls, layer, flags, qb, qa, la, sw,
                ops, il, op, fq, js)
        finally:
            self.lib.qref_result_destroy(h)

    def map_qasm(self, arch: "RefArch", qasm_file: str, cfg: QmgConfig, log_path: str = "") -> RefMapResult:
        h = self.lib.qref_map_qasm(arch.h, qasm_file.encode(), C.byref(cfg), log_path.encode())
        return self._collect(arch, h)

    def map_gates(self, arch: "RefArch", n_qubits, control, target, cfg: QmgConfig, log_path: str = "") -> RefMapResult:
        c = np.ascontiguousarray(control, dtype=np.int32)
        t = np.ascontiguousarray(target, dtype=np.int32)
        h = self.lib.qref_map_gates(arch.h, n_qubits, len(c), c.ctypes.data, t.ctypes.data,
                                    C.byref(cfg), log_path.encode())
        return self._collect(arch, h)


class RefArch:
    def __init__(self, ref: Ref, h):
        self.ref, self.h = ref, h
        info = QmgArchInfo()
        ref.lib.qref_arch_info(h, C.byref(info))
        self.info = info
        self.n = int(info.n)
        self.n_edges = int(info.n_edges)

    def __del__(self):
        if getattr(self, "h", None):
            self.ref.lib.qref_arch_destroy(self.h)
            self.h = None

    def refresh(self):
        self.ref.lib.qref_arch_info(self.h, C.byref(self.info))

    @property
    def edges(self) -> np.ndarray:
        e = np.zeros((self.n_edges, 2), dtype=np.uint16)
        self.ref.lib.qref_arch_edges(self.h, e.ctypes.data)
        return e

    def set_error_rates(self, q1err, cxerr):
        a = np.ascontiguousarray(q1err, dtype=np.float64)
        b = np.ascontiguousarray(cxerr, dtype=np.float64)
        assert len(a) == self.n and len(b) == self.n_edges
        if self.ref.lib.qref_arch_set_error_rates(self.h, a.ctypes.data, b.ctypes.data):
            raise RuntimeError(self.ref._err())
        self.refresh()

    def minimum_number_of_swaps(self, permutation, limit=-1) -> int:
        L = self.ref.lib
        L.qref_arch_minimum_number_of_swaps.restype = C.c_longlong
        L.qref_arch_minimum_number_of_swaps.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_longlong]
        p = np.ascontiguousarray(permutation, dtype=np.uint16)
        r = L.qref_arch_minimum_number_of_swaps(self.h, p.ctypes.data, len(p), int(limit))
        if r == -2:
            raise RuntimeError(self.ref._err())
        return int(r)

    def coupling_limit(self, qubit_choice=None) -> int:
        L = self.ref.lib
        L.qref_arch_coupling_limit.restype = C.c_longlong
        L.qref_arch_coupling_limit.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        if qubit_choice is None:
            return int(L.qref_arch_coupling_limit(self.h, None, 0))
        c = np.ascontiguousarray(sorted(qubit_choice), dtype=np.uint16)
        return int(L.qref_arch_coupling_limit(self.h, c.ctypes.data, len(c)))

    def distance_table(self, include_reversal_cost=True) -> np.ndarray:
        out = np.empty((self.n, self.n), dtype=np.float64)
        if self.ref.lib.qref_arch_distance_table(self.h, int(include_reversal_cost), out.ctypes.data):
            raise RuntimeError(self.ref._err())
        return out

    def _fid(self, which, shape):
        out = np.empty(shape, dtype=np.float64)
        if self.ref.lib.qref_arch_fidelity(self.h, which, out.ctypes.data):
            raise RuntimeError(self.ref._err())
        return out

    def single_qubit_fidelity_costs(self):
        return self._fid(0, (self.n,))

    def two_qubit_fidelity_costs(self):
        return self._fid(1, (self.n, self.n))

    def swap_fidelity_costs(self):
        return self._fid(2, (self.n, self.n))

    def fidelity_distance_tables(self):
        self.refresh()
        return self._fid(3, (int(self.info.n_fidelity_tables), self.n, self.n))
