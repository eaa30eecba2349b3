"""Batch routing front-end: packs routing instances into the C-ABI structs of
include/qmap_b200.h, owns the host result buffers and unpacks them.

The compute call itself is `qmg_route_batch` (CUDA); the packing code is
written against a small "backend" protocol so that tests can point it at the
host debug build of the kernel source (tests/emu) where no GPU exists.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import _abi
from ._abi import (LAYER_RESULT_DTYPE, NODE_LOG_DTYPE, QmgConfig, QmgInstance, QmgLayers,
                   QmgRouteOptions, QmgRouteResult, QmgRouteStats, QmapError)


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
    debug=False,
    verbose=False,
) -> QmgConfig:
    """Configuration with the reference's defaults
    (include/sc/configuration/Configuration.hpp:30-119)."""

    def enum(table, v):
        if isinstance(v, str):
            if v not in table:
                raise ValueError(f"Invalid value: {v}")
            return table[v]
        return int(v)

    c = QmgConfig()
    c.heuristic = enum(_abi.HEURISTICS, heuristic)
    c.lookahead_heuristic = enum(_abi.LOOKAHEADS, lookahead_heuristic)
    c.nr_lookaheads = nr_lookaheads
    c.first_lookahead_factor = first_lookahead_factor
    c.lookahead_factor = lookahead_factor
    c.early_termination = enum(_abi.EARLY_TERMINATIONS, early_termination)
    c.early_termination_limit = early_termination_limit
    c.automatic_layer_splits = int(automatic_layer_splits)
    c.automatic_layer_splits_node_limit = automatic_layer_splits_node_limit
    c.layering = enum(_abi.LAYERINGS, layering)
    c.initial_layout = enum(_abi.LAYOUTS, initial_layout)
    c.iterative_bidirectional_routing_passes = iterative_bidirectional_routing_passes
    c.swap_on_first_layer = int(swap_on_first_layer)
    c.add_barriers_between_layers = int(add_barriers_between_layers)
    c.add_measurements_to_mapped_circuit = int(add_measurements_to_mapped_circuit)
    c.pre_mapping_optimizations = int(pre_mapping_optimizations)
    c.post_mapping_optimizations = int(post_mapping_optimizations)
    c.debug = int(debug)
    c.verbose = int(verbose)
    return c


class LayeredCircuit:
    """qmg_circuit: Mapper::createLayers on a gate list over logical qubits
    (control < 0 marks a single-qubit gate)."""

    def __init__(self, lib, n_arch_qubits: int, control, target, layering=0):
        self.lib = lib
        self.n = int(n_arch_qubits)
        self.control = np.ascontiguousarray(control, dtype=np.int32)
        self.target = np.ascontiguousarray(target, dtype=np.int32)
        assert self.control.shape == self.target.shape
        h = C.c_void_p()
        _abi.check(lib, lib.qmg_circuit_create(
            self.n, len(self.control), self.control.ctypes.data,
            self.target.ctypes.data, int(layering), C.byref(h)))
        self.h = h
        self.layering = int(layering)
        self.splits = []  # layers split so far, in order

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.qmg_circuit_destroy(self.h)
            self.h = None

    def split_copy(self, layer) -> "LayeredCircuit":
        """A new circuit with the same gates and split history plus a split of
        `layer` (Mapper::splitLayer; a sequence = several splits in that order,
        e.g. RouteOutput.split_layers); this object is left untouched, other
        routing instances may share it."""
        out = LayeredCircuit(self.lib, self.n, self.control, self.target, self.layering)
        more = [int(layer)] if np.isscalar(layer) else [int(l) for l in layer]
        for l in self.splits + more:
            out.split(l)
        return out

    @property
    def layers(self) -> QmgLayers:
        v = QmgLayers()
        self.lib.qmg_circuit_layers(self.h, C.byref(v))
        return v

    @property
    def n_layers(self) -> int:
        return int(self.layers.n_layers)

    @property
    def n_2q_gates(self) -> int:
        return int((self.control >= 0).sum())

    def layer_gates(self):
        """(offsets, gate indices): the gates of every layer in layer order."""
        nl = self.n_layers
        off = np.zeros(nl + 1, dtype=np.uint32)
        idx = np.zeros(len(self.control), dtype=np.uint32)
        self.lib.qmg_circuit_layer_gates(self.h, off.ctypes.data, idx.ctypes.data)
        return off, idx

    def splittable(self, layer: int) -> bool:
        return bool(self.lib.qmg_circuit_layer_splittable(self.h, layer))

    def split(self, layer: int):
        _abi.check(self.lib, self.lib.qmg_circuit_split_layer(self.h, layer))
        self.splits.append(int(layer))


class CircuitBatch:
    """qmg_circuit_batch: Mapper::createLayers for many gate lists at once (host
    threads, CSR layers only).  batch[i] is a light view usable as
    RouteItem.circuit."""

    class View:
        def __init__(self, batch, i):
            self.batch, self.i = batch, i

        @property
        def layers(self) -> QmgLayers:
            v = QmgLayers()
            _abi.check(self.batch.lib, self.batch.lib.qmg_circuit_batch_layers(
                self.batch.h, self.i, C.byref(v)))
            return v

        @property
        def n_layers(self) -> int:
            return int(self.layers.n_layers)

        def layer_gates(self):
            lo, hi = self.batch.offsets[self.i], self.batch.offsets[self.i + 1]
            off = np.zeros(self.n_layers + 1, dtype=np.uint32)
            idx = np.zeros(int(hi - lo), dtype=np.uint32)
            _abi.check(self.batch.lib, self.batch.lib.qmg_circuit_batch_layer_gates(
                self.batch.h, self.i, off.ctypes.data, idx.ctypes.data))
            return off, idx

    def __init__(self, lib, n_arch_qubits: int, offsets, control, target, layering=0):
        self.lib = lib
        self.n = int(n_arch_qubits)
        self.offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        self.control = np.ascontiguousarray(control, dtype=np.int32)
        self.target = np.ascontiguousarray(target, dtype=np.int32)
        h = C.c_void_p()
        _abi.check(lib, lib.qmg_circuit_batch_create(
            self.n, len(self.offsets) - 1, self.offsets.ctypes.data, self.control.ctypes.data,
            self.target.ctypes.data, int(layering), C.byref(h)))
        self.h = h

    @classmethod
    def from_gate_lists(cls, lib, n_arch_qubits, gate_lists, layering=0):
        """gate_lists: [(control, target)] arrays per circuit"""
        sizes = [len(c) for c, _ in gate_lists]
        off = np.concatenate(([0], np.cumsum(sizes))).astype(np.uint64)
        ctrl = np.concatenate([np.asarray(c, dtype=np.int32) for c, _ in gate_lists]) \
            if gate_lists else np.zeros(0, np.int32)
        tgt = np.concatenate([np.asarray(t, dtype=np.int32) for _, t in gate_lists]) \
            if gate_lists else np.zeros(0, np.int32)
        return cls(lib, n_arch_qubits, off, ctrl, tgt, layering)

    @classmethod
    def from_qasm_files(cls, lib, n_arch_qubits, paths, layering=0):
        """Files -> layers without a Python object per gate: the library's
        OpenQASM 2 reader (qmg_qasm_*), the mapper's view of each circuit
        (idle qubits stripped) and createLayers of the whole batch.  Returns
        (batch, [identity initial locations per circuit])."""
        from . import circuit as qcirc
        views = [qcirc.load_qasm_native(lib, p).mapper_view(n_arch_qubits) for p in paths]
        batch = cls.from_gate_lists(lib, n_arch_qubits, [(c, t) for c, t, _ in views], layering)
        return batch, [ident for _, _, ident in views]

    def __len__(self):
        return len(self.offsets) - 1

    def __getitem__(self, i):
        return CircuitBatch.View(self, int(i))

    def fill_instances(self, inst_array, n):
        _abi.check(self.lib, self.lib.qmg_circuit_batch_fill_instances(
            self.h, C.addressof(inst_array), n))

    def verify(self, edges, res_array, n_instances, fidelity_aware=False) -> np.ndarray:
        """qmg_circuit_batch_verify: replay of the routed instances; returns the
        verdicts (1 valid, 0 invalid, 2 not checkable)."""
        e = np.ascontiguousarray(edges, dtype=np.uint16).reshape(-1, 2)
        v = np.zeros(n_instances, dtype=np.uint8)
        _abi.check(self.lib, self.lib.qmg_circuit_batch_verify(
            self.h, self.n, e.ctypes.data, len(e), self.offsets.ctypes.data,
            self.control.ctypes.data, self.target.ctypes.data, C.addressof(res_array),
            n_instances, int(bool(fidelity_aware)), v.ctypes.data))
        return v

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.qmg_circuit_batch_destroy(self.h)
            self.h = None


@dataclass
class RouteItem:
    circuit: LayeredCircuit
    initial_locations: np.ndarray | None = None  # int16[n], -1 = unplaced
    bidirectional_passes: int = 0
    first_layer: int = 0
    n_route_layers: int = 0


@dataclass
class RouteOutput:
    status: int
    n_layers_done: int
    expanded_nodes: int
    generated_nodes: int
    layers: np.ndarray  # LAYER_RESULT_DTYPE per layer
    swaps: np.ndarray  # (n_swaps_total, 2) uint16
    placements: np.ndarray  # (n_placements, 2) uint16 (logical, physical)
    final_qubits: np.ndarray
    final_locations: np.ndarray
    start_locations: np.ndarray
    layer_qubits: np.ndarray | None = None
    layer_n_swaps: np.ndarray | None = None  # uint16 per layer
    split_layers: np.ndarray | None = None  # layers split by the backend (final pass)
    node_log: np.ndarray | None = None  # NODE_LOG_DTYPE, generation order (final pass)
    n_node_log: int = 0  # nodes generated (more than len(node_log): the log was cut)
    layer_root_qubits: np.ndarray | None = None  # (layers, n): root layout of every layer
    expanded_nodes_final_pass: int = 0  # the reference's heuristicBenchmark counters
    generated_nodes_final_pass: int = 0
    _offsets: np.ndarray | None = None

    def layer_swaps(self, layer: int) -> np.ndarray:
        if self.layers is not None:
            r = self.layers[layer]
            return self.swaps[r["swap_offset"]: r["swap_offset"] + r["n_swaps"]]
        if self._offsets is None:
            self._offsets = np.concatenate(([0], np.cumsum(self.layer_n_swaps, dtype=np.int64)))
        return self.swaps[self._offsets[layer]: self._offsets[layer + 1]]


@dataclass
class BatchBuffers:
    """Host result buffers of a batch (kept alive while the C side uses them)."""
    n: int
    items: list
    inst: C.Array = None
    res: C.Array = None
    layers: list = field(default_factory=list)
    swaps: list = field(default_factory=list)
    places: list = field(default_factory=list)
    fq: np.ndarray = None
    fl: np.ndarray = None
    sl: np.ndarray = None
    lq: list = field(default_factory=list)
    lns: list = field(default_factory=list)
    keep: list = field(default_factory=list)
    layer_cap: list = field(default_factory=list)  # per-instance rows of layers/lq/lns
    node_logs: list = field(default_factory=list)
    root_q: list = field(default_factory=list)
    split_logs: list = field(default_factory=list)


def pack(n: int, items, want_layer_qubits=False, swap_capacity=None, lean=False,
         extra_layers=0, node_log_capacity=0) -> BatchBuffers:
    """lean: no per-layer qmg_layer_result records, only the compact per-layer
    SWAP counts (what circuit emission needs).  extra_layers: additional rows
    in the per-layer outputs; with extra_layers > 0 the backend splits layers
    itself (qmg_instance.layer_capacity) and reports the splits of the final
    pass in RouteOutput.split_layers."""
    nI = len(items)
    b = BatchBuffers(n=n, items=list(items))
    b.inst = (QmgInstance * nI)()
    b.res = (QmgRouteResult * nI)()
    b.fq = np.full((nI, n), -1, dtype=np.int16)
    b.fl = np.full((nI, n), -1, dtype=np.int16)
    b.sl = np.full((nI, n), -1, dtype=np.int16)
    for i, it in enumerate(items):
        lay = it.circuit.layers
        nl = int(lay.n_layers)
        cap_l = max(nl + extra_layers, 1)
        b.layer_cap.append(cap_l)
        ins = b.inst[i]
        ins.layers = lay
        if it.initial_locations is not None:
            loc = np.ascontiguousarray(it.initial_locations, dtype=np.int16)
            assert loc.shape == (n,)
            b.keep.append(loc)
            ins.initial_locations = loc.ctypes.data
        ins.bidirectional_passes = it.bidirectional_passes
        ins.first_layer = it.first_layer
        ins.n_route_layers = it.n_route_layers
        if extra_layers > 0:
            ins.layer_capacity = cap_l
            slog = np.zeros(extra_layers, dtype=np.uint32)
            b.split_logs.append(slog)
            b.res[i].split_layers = slog.ctypes.data
            b.res[i].split_capacity = extra_layers
        cap = swap_capacity if swap_capacity is not None else 4 * cap_l + 256
        lr = None if lean else np.zeros(cap_l, dtype=LAYER_RESULT_DTYPE)
        sw = np.zeros((max(cap, 1), 2), dtype=np.uint16)
        pl = np.zeros((n, 2), dtype=np.uint16)
        b.layers.append(lr)
        b.swaps.append(sw)
        b.places.append(pl)
        r = b.res[i]
        if lr is not None:
            r.layers = lr.ctypes.data
        else:
            ln = np.zeros(cap_l, dtype=np.uint16)
            b.lns.append(ln)
            r.layer_n_swaps = ln.ctypes.data
        r.swaps = sw.ctypes.data
        r.swap_capacity = cap
        r.placements = pl.ctypes.data
        r.placement_capacity = n
        r.final_qubits = b.fq[i].ctypes.data
        r.final_locations = b.fl[i].ctypes.data
        r.start_locations = b.sl[i].ctypes.data
        if node_log_capacity > 0:
            # DataLogger-style log of every generated node (final pass)
            nlog = np.zeros(node_log_capacity, dtype=NODE_LOG_DTYPE)
            rq = np.full((cap_l, n), -1, dtype=np.int16)
            b.node_logs.append(nlog)
            b.root_q.append(rq)
            r.node_log = nlog.ctypes.data
            r.node_log_capacity = node_log_capacity
            r.layer_root_qubits = rq.ctypes.data
        if want_layer_qubits:
            lq = np.full((cap_l, n), -1, dtype=np.int16)
            b.lq.append(lq)
            r.layer_qubits = lq.ctypes.data
    return b


def unpack(b: BatchBuffers) -> list:
    outs = []
    for i, it in enumerate(b.items):
        r = b.res[i]
        nl = min(max(it.circuit.n_layers, int(r.n_layers_done)), b.layer_cap[i])
        outs.append(RouteOutput(
            status=int(r.status), n_layers_done=int(r.n_layers_done),
            expanded_nodes=int(r.expanded_nodes), generated_nodes=int(r.generated_nodes),
            layers=b.layers[i][:nl] if b.layers[i] is not None else None,
            layer_n_swaps=b.lns[i][:nl] if b.lns else None,
            swaps=b.swaps[i][: min(int(r.n_swaps_total), len(b.swaps[i]))],
            placements=b.places[i][: int(r.n_placements_total)],
            final_qubits=b.fq[i], final_locations=b.fl[i], start_locations=b.sl[i],
            layer_qubits=b.lq[i][:nl] if b.lq else None,
            node_log=(b.node_logs[i][: min(int(r.n_node_log), len(b.node_logs[i]))]
                      if b.node_logs else None),
            n_node_log=int(r.n_node_log),
            layer_root_qubits=b.root_q[i][:nl] if b.root_q else None,
            expanded_nodes_final_pass=int(r.expanded_nodes_final_pass),
            generated_nodes_final_pass=int(r.generated_nodes_final_pass),
            split_layers=(b.split_logs[i][: min(int(r.n_splits), len(b.split_logs[i]))]
                          if b.split_logs else None)))
    return outs


class CudaBackend:
    """qmg_route_batch on a device-resident Architecture handle."""

    def __init__(self, lib, arch_handle):
        self.lib, self.arch = lib, arch_handle
        self.stats = QmgRouteStats()

    def route(self, cfg: QmgConfig, b: BatchBuffers, options: QmgRouteOptions | None = None):
        rc = self.lib.qmg_route_batch(
            self.arch, C.byref(cfg), len(b.items), C.addressof(b.inst),
            C.addressof(b.res), C.byref(options) if options is not None else None,
            C.byref(self.stats))
        _abi.check(self.lib, rc)


def route_batch(backend, n: int, cfg: QmgConfig, items, want_layer_qubits=False,
                options=None, max_retries=6, lean=False, extra_layers=0,
                node_log_capacity=0) -> list:
    """Route `items` (RouteItem) and return one RouteOutput per item.  Swap
    buffers that turn out too small are grown and the batch is re-run."""
    cap = None
    for _ in range(max_retries):
        b = pack(n, items, want_layer_qubits, cap, lean, extra_layers, node_log_capacity)
        if options is not None:
            backend.route(cfg, b, options)
        else:
            backend.route(cfg, b)
        outs = unpack(b)
        if not any(o.status == _abi.QMG_BUFFER_TOO_SMALL for o in outs):
            return outs
        cur = max(int(r.swap_capacity) for r in b.res)
        cap = cur * 4
    raise QmapError("swap buffer capacity could not be satisfied")
