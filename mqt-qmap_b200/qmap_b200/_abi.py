"""ctypes mirror of include/qmap_b200.h and the loader of libqmap_b200.so.

There is no CPU fallback: if the CUDA library is missing `load()` raises, and
without a CUDA device every compute entry point returns QMG_NO_DEVICE.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get(
    "QMAP_B200_LIB", os.path.join(os.path.dirname(HERE), "csrc", "libqmap_b200.so"))

QMG_OK, QMG_NO_VIABLE_MAPPING, QMG_SPLIT_LIMIT_HIT, QMG_BAD_ARG = 0, 1, 2, 3
QMG_CUDA_ERROR, QMG_OOM, QMG_POOL_EXHAUSTED, QMG_NO_DEVICE = 4, 5, 6, 7
QMG_BUFFER_TOO_SMALL = 8
STATUS_NAMES = {
    0: "ok", 1: "No viable mapping found.", 2: "layer split limit hit",
    3: "bad argument", 4: "CUDA error", 5: "out of memory",
    6: "node pool exhausted", 7: "no CUDA device", 8: "output buffer too small",
}

HEURISTICS = {
    "gate_count_max_distance": 0,
    "gate_count_sum_distance": 1,
    "gate_count_sum_distance_minus_shared_swaps": 2,
    "gate_count_max_distance_or_sum_distance_minus_shared_swaps": 3,
    "fidelity_best_location": 4,
}
LOOKAHEADS = {"none": 0, "gate_count_max_distance": 1, "gate_count_sum_distance": 2}
LAYERINGS = {"individual_gates": 0, "disjoint_qubits": 1, "odd_gates": 2,
             "qubit_triangle": 3, "disjoint_2q_blocks": 4}
LAYOUTS = {"identity": 0, "static": 1, "dynamic": 2}
EARLY_TERMINATIONS = {
    "none": 0, "expanded_nodes": 1, "expanded_nodes_after_first_solution": 2,
    "expanded_nodes_after_current_optimal_solution": 3, "solution_nodes": 4,
    "solution_nodes_after_current_optimal_solution": 5,
}


class QmgConfig(C.Structure):
    _fields_ = [
        ("heuristic", C.c_int32), ("lookahead_heuristic", C.c_int32),
        ("nr_lookaheads", C.c_uint32), ("_pad0", C.c_uint32),
        ("first_lookahead_factor", C.c_double), ("lookahead_factor", C.c_double),
        ("early_termination", C.c_int32), ("automatic_layer_splits", C.c_int32),
        ("early_termination_limit", C.c_uint64),
        ("automatic_layer_splits_node_limit", C.c_uint64),
        ("layering", C.c_int32), ("initial_layout", C.c_int32),
        ("iterative_bidirectional_routing_passes", C.c_uint32),
        ("swap_on_first_layer", C.c_int32), ("add_barriers_between_layers", C.c_int32),
        ("add_measurements_to_mapped_circuit", C.c_int32),
        ("pre_mapping_optimizations", C.c_int32), ("post_mapping_optimizations", C.c_int32),
        ("debug", C.c_int32), ("verbose", C.c_int32),
    ]


class QmgLayers(C.Structure):
    _fields_ = [
        ("n_layers", C.c_uint32), ("n_logical", C.c_uint32),
        ("gate2q_offsets", C.c_void_p), ("gates2q", C.c_void_p),
        ("gate1q_offsets", C.c_void_p), ("gates1q", C.c_void_p),
    ]


class QmgInstance(C.Structure):
    _fields_ = [
        ("layers", QmgLayers), ("initial_locations", C.c_void_p),
        ("bidirectional_passes", C.c_uint32), ("first_layer", C.c_uint32),
        ("n_route_layers", C.c_uint32), ("layer_capacity", C.c_uint32),
    ]


class QmgRouteResult(C.Structure):
    _fields_ = [
        ("layers", C.c_void_p), ("swaps", C.c_void_p),
        ("swap_capacity", C.c_uint32), ("n_swaps_total", C.c_uint32),
        ("placements", C.c_void_p),
        ("placement_capacity", C.c_uint32), ("n_placements_total", C.c_uint32),
        ("final_qubits", C.c_void_p), ("final_locations", C.c_void_p),
        ("start_locations", C.c_void_p), ("layer_qubits", C.c_void_p),
        ("n_layers_done", C.c_uint32), ("status", C.c_int32),
        ("expanded_nodes", C.c_uint64), ("generated_nodes", C.c_uint64),
        ("layer_n_swaps", C.c_void_p),
        ("split_layers", C.c_void_p), ("split_capacity", C.c_uint32),
        ("n_splits", C.c_uint32),
        ("expanded_nodes_final_pass", C.c_uint64), ("generated_nodes_final_pass", C.c_uint64),
        ("node_log", C.c_void_p), ("node_log_capacity", C.c_uint32), ("n_node_log", C.c_uint32),
        ("layer_root_qubits", C.c_void_p),
        ("placement_log", C.c_void_p), ("placement_log_capacity", C.c_uint32),
        ("n_placement_log", C.c_uint32),
    ]


class QmgRouteOptions(C.Structure):
    _fields_ = [("pool_nodes", C.c_uint32), ("threads_per_search", C.c_uint32),
                ("searches_per_sm", C.c_uint32), ("lean_outputs", C.c_uint32),
                ("max_searches", C.c_uint32), ("search_variant", C.c_uint32),
                ("big_pools", C.c_uint32), ("big_pool_nodes", C.c_uint32)]


class QmgRouteStats(C.Structure):
    _fields_ = [
        ("kernel_ms", C.c_float), ("h2d_ms", C.c_float), ("d2h_ms", C.c_float),
        ("grid", C.c_uint32), ("block", C.c_uint32),
        ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64),
        ("expanded_nodes", C.c_uint64), ("generated_nodes", C.c_uint64),
        ("kernel_launches", C.c_uint32), ("retried_instances", C.c_uint32),
        ("big_pool_layers", C.c_uint32),
    ]


class QmgArchInfo(C.Structure):
    _fields_ = [
        ("n", C.c_uint16), ("_pad", C.c_uint16), ("n_edges", C.c_uint32),
        ("bidirectional", C.c_int32), ("unidirectional", C.c_int32),
        ("fidelity_available", C.c_int32), ("n_fidelity_tables", C.c_uint32),
        ("device", C.c_int32), ("table_build_ms", C.c_float),
    ]


LAYER_RESULT_DTYPE = np.dtype([
    ("cost_fixed", "f8"), ("cost_fixed_reversals", "f8"), ("cost_heur", "f8"),
    ("lookahead_penalty", "f8"), ("expanded_nodes", "u8"), ("generated_nodes", "u8"),
    ("expanded_nodes_after_first_solution", "u8"),
    ("expanded_nodes_after_optimal_solution", "u8"), ("solution_nodes", "u8"),
    ("solution_nodes_after_optimal_solution", "u8"), ("layout_hash", "u8"),
    ("solution_depth", "u4"), ("result_node_id", "u4"), ("swap_offset", "u4"),
    ("n_swaps", "u4"), ("placement_offset", "u4"), ("n_placements", "u4"),
    ("early_termination", "i4"), ("status", "i4"),
])
assert LAYER_RESULT_DTYPE.itemsize == 120
NODE_LOG_DTYPE = np.dtype([
    ("cost_fixed", "f8"), ("cost_heur", "f8"), ("lookahead_penalty", "f8"),
    ("id", "u4"), ("parent", "u4"), ("depth", "u4"), ("layer", "u4"),
    ("swap_first", "u2"), ("swap_second", "u2"), ("valid", "u2"), ("root", "u2"),
])
assert NODE_LOG_DTYPE.itemsize == 48

# every symbol include/qmap_b200.h declares
EXPORTS = [
    "qmg_config_default", "qmg_build_table", "qmg_build_single_edge_skip_table",
    "qmg_build_edge_skip_table", "qmg_arch_create", "qmg_arch_set_fidelity_costs",
    "qmg_arch_destroy", "qmg_arch_get_info", "qmg_arch_get_distance_table",
    "qmg_arch_get_fidelity_distance_tables", "qmg_circuit_create",
    "qmg_circuit_layers", "qmg_circuit_layer_gates", "qmg_circuit_layer_splittable",
    "qmg_circuit_split_layer", "qmg_circuit_destroy", "qmg_route_batch",
    "qmg_batch_create", "qmg_batch_run", "qmg_batch_fetch", "qmg_batch_destroy",
    "qmg_last_error", "qmg_device_count", "qmg_version", "qmg_build_table_static",
    "qmg_circuit_batch_create", "qmg_circuit_batch_layers", "qmg_circuit_batch_destroy",
    "qmg_circuit_batch_fill_instances", "qmg_circuit_batch_layer_gates",
    "qmg_circuit_batch_verify",
    "qmg_qasm_read_file", "qmg_qasm_parse", "qmg_qasm_get_info", "qmg_qasm_gates",
    "qmg_qasm_measurements", "qmg_qasm_mapper_view", "qmg_qasm_gate_name", "qmg_qasm_destroy",
]

_lib = None


class QmgQasmInfo(C.Structure):
    _fields_ = [("n_qubits", C.c_uint32), ("n_clbits", C.c_uint32),
                ("n_gates", C.c_uint64), ("n_measurements", C.c_uint64)]


def bind_host_api(L):
    """argtypes of the host-only entry points (shared with the test emulator)."""
    L.qmg_last_error.restype = C.c_char_p
    L.qmg_config_default.argtypes = [C.POINTER(QmgConfig)]
    L.qmg_circuit_create.argtypes = [C.c_uint32, C.c_size_t, C.c_void_p, C.c_void_p,
                                     C.c_int32, C.POINTER(C.c_void_p)]
    L.qmg_circuit_layers.argtypes = [C.c_void_p, C.POINTER(QmgLayers)]
    L.qmg_circuit_layer_gates.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.qmg_circuit_layer_splittable.argtypes = [C.c_void_p, C.c_uint32]
    L.qmg_circuit_split_layer.argtypes = [C.c_void_p, C.c_uint32]
    L.qmg_circuit_destroy.argtypes = [C.c_void_p]
    L.qmg_circuit_destroy.restype = None
    L.qmg_circuit_batch_create.argtypes = [C.c_uint32, C.c_size_t, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_int32, C.POINTER(C.c_void_p)]
    L.qmg_circuit_batch_layers.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(QmgLayers)]
    L.qmg_circuit_batch_fill_instances.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
    L.qmg_circuit_batch_layer_gates.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
    L.qmg_circuit_batch_verify.argtypes = [C.c_void_p, C.c_uint16, C.c_void_p, C.c_size_t, C.c_void_p,
                                           C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int32,
                                           C.c_void_p]
    L.qmg_circuit_batch_destroy.argtypes = [C.c_void_p]
    L.qmg_circuit_batch_destroy.restype = None
    L.qmg_qasm_read_file.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
    L.qmg_qasm_parse.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(C.c_void_p)]
    L.qmg_qasm_get_info.argtypes = [C.c_void_p, C.POINTER(QmgQasmInfo)]
    L.qmg_qasm_gates.argtypes = [C.c_void_p] * 6
    L.qmg_qasm_measurements.argtypes = [C.c_void_p] * 3
    L.qmg_qasm_mapper_view.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
    L.qmg_qasm_gate_name.argtypes = [C.c_uint8]
    L.qmg_qasm_gate_name.restype = C.c_char_p
    L.qmg_qasm_destroy.argtypes = [C.c_void_p]
    L.qmg_qasm_destroy.restype = None


def load(path: str = LIB_PATH):
    """Load libqmap_b200.so (built by __graft_entry__.build()); fails loudly."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(path):
        raise RuntimeError(
            f"{path} is missing: the CUDA library has not been built "
            "(run `python -c 'import __graft_entry__ as g; g.build()'`). "
            "There is no CPU fallback.")
    L = C.CDLL(path)
    bind_host_api(L)
    vp, u16, sz, dbl, i32 = C.c_void_p, C.c_uint16, C.c_size_t, C.c_double, C.c_int
    L.qmg_version.restype = C.c_char_p
    L.qmg_device_count.restype = C.c_int
    L.qmg_build_table.argtypes = [u16, vp, sz, vp, i32, vp, i32]
    L.qmg_build_table_static.argtypes = [u16, vp, sz, vp, vp, i32, vp]
    L.qmg_build_single_edge_skip_table.argtypes = [u16, vp, vp, sz, dbl, vp, i32]
    L.qmg_build_edge_skip_table.argtypes = [u16, vp, sz, vp, i32, vp, sz,
                                            C.POINTER(C.c_size_t), i32]
    L.qmg_arch_create.argtypes = [u16, vp, sz, i32, C.POINTER(vp)]
    L.qmg_arch_set_fidelity_costs.argtypes = [vp, vp, vp, vp]
    L.qmg_arch_destroy.argtypes = [vp]
    L.qmg_arch_destroy.restype = None
    L.qmg_arch_get_info.argtypes = [vp, C.POINTER(QmgArchInfo)]
    L.qmg_arch_get_distance_table.argtypes = [vp, i32, vp]
    L.qmg_arch_get_fidelity_distance_tables.argtypes = [vp, vp]
    L.qmg_route_batch.argtypes = [vp, C.POINTER(QmgConfig), sz, vp, vp,
                                  C.POINTER(QmgRouteOptions), C.POINTER(QmgRouteStats)]
    L.qmg_batch_create.argtypes = [vp, C.POINTER(QmgConfig), sz, vp,
                                   C.POINTER(QmgRouteOptions), C.POINTER(vp)]
    L.qmg_batch_run.argtypes = [vp, C.POINTER(QmgRouteStats)]
    L.qmg_batch_fetch.argtypes = [vp, vp]
    L.qmg_batch_destroy.argtypes = [vp]
    L.qmg_batch_destroy.restype = None
    _lib = L
    return L


class QmapError(RuntimeError):
    """Mirror of QMAPException (include/sc/utils.hpp:46-56)."""


def check(L, rc: int):
    if rc != QMG_OK:
        msg = L.qmg_last_error().decode() or STATUS_NAMES.get(rc, str(rc))
        raise QmapError(msg)
