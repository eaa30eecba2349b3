"""Host mirror of the reference's HeuristicMapper: the same call
(`HeuristicMapper(circuit, architecture).map(config)`), with the per-layer A*
searches done by the CUDA library (one launch routes the whole circuit).

What stays on the host is what the reference does around `aStarMap`
(src/sc/heuristic/HeuristicMapper.cpp): parameter checks (:149-167), the
initial layout (:109-195), the layer-split-and-resume loop, and the emission of
the mapped circuit with its counters (`routeCircuit`, :346-519)."""
from __future__ import annotations

import time
from dataclasses import dataclass, field

import numpy as np

from . import _abi, routing
from ._abi import QmapError
from .circuit import Circuit, Op


class InvalidPlacement(QmapError):
    """a routed layer leaves a two-qubit gate on qubits that are not coupled"""


@dataclass
class MappingResults:
    """The fields of MappingResults (include/sc/MappingResults.hpp) this path
    produces."""
    input_gates: int = 0
    input_cnots: int = 0
    input_single_qubit_gates: int = 0
    input_layers: int = 0
    output_gates: int = 0
    output_swaps: int = 0
    output_direction_reverse: int = 0
    output_cnots: int = 0
    output_single_qubit_gates: int = 0
    layer_splits: int = 0
    rerouted_with_host_splits: bool = False
    expanded_nodes: int = 0  # final routing pass (heuristicBenchmark.expandedNodes)
    generated_nodes: int = 0
    expanded_nodes_all_passes: int = 0  # incl. pseudo-routing passes
    generated_nodes_all_passes: int = 0
    time: float = 0.0
    layer_n_swaps: np.ndarray | None = None
    config: object = None
    extras: dict = field(default_factory=dict)
    # CircuitInfo / statistics as MappingResults::json() reports them
    input_name: str = ""
    input_qubits: int = 0
    output_name: str = ""
    output_qubits: int = 0
    architecture: str = ""
    timeout: bool = True
    total_fidelity: float = 1.0
    total_log_fidelity: float = 0.0
    # heuristicBenchmark (filled with config.debug, HeuristicMapper.cpp:432-444)
    average_branching_factor: float = 0.0
    effective_branching_factor: float = 0.0
    seconds_per_node: float = 0.0
    layer_benchmark: np.ndarray | None = None  # per-layer records (debug)

    def json(self) -> dict:
        """MappingResults::json() (include/sc/MappingResults.hpp:127-175) with
        Configuration::json() (src/sc/Configuration.cpp:24-78)."""
        cfg = self.config

        def name(table, v):
            return next(k for k, x in table.items() if x == v)

        heur = name(_abi.HEURISTICS, cfg.heuristic)
        fid = heur == "fidelity_best_location"
        settings = {
            "heuristic": heur,
            # isAdmissible / isPrincipallyAdmissible / isTight / isFidelityAware
            # (include/sc/configuration/Heuristic.hpp:44-126)
            "heuristic_properties": {
                "admissible": heur in ("gate_count_max_distance", "fidelity_best_location"),
                "principally_admissible": heur != "gate_count_sum_distance",
                "tight": not fid,
                "fidelity_aware": fid,
            },
            "initial_layout": name(_abi.LAYOUTS, cfg.initial_layout),
        }
        if cfg.lookahead_heuristic != _abi.LOOKAHEADS["none"]:
            settings["lookahead"] = {
                "heuristic": name(_abi.LOOKAHEADS, cfg.lookahead_heuristic),
                "heuristic_properties": {"fidelity_aware": False},
                "lookaheads": int(cfg.nr_lookaheads),
                "first_factor": float(cfg.first_lookahead_factor),
                "factor": float(cfg.lookahead_factor),
            }
        return {
            "circuit": {"name": self.input_name, "qubits": self.input_qubits, "gates": self.input_gates,
                        "single_qubit_gates": self.input_single_qubit_gates, "cnots": self.input_cnots},
            "mapped_circuit": {"name": self.output_name, "qubits": self.output_qubits,
                               "gates": self.output_gates,
                               "single_qubit_gates": self.output_single_qubit_gates,
                               "cnots": self.output_cnots},
            "config": {
                "method": "heuristic",
                "layering_strategy": name(_abi.LAYERINGS, cfg.layering),
                "pre_mapping_optimizations": bool(cfg.pre_mapping_optimizations),
                "post_mapping_optimizations": bool(cfg.post_mapping_optimizations),
                "add_measurements_to_mapped_circuit": bool(cfg.add_measurements_to_mapped_circuit),
                "verbose": bool(cfg.verbose),
                "debug": bool(cfg.debug),
                "settings": settings,
            },
            "statistics": {
                "timeout": self.timeout,
                "mapping_time": self.time,
                "arch": self.architecture,
                "layers": self.input_layers,
                "swaps": self.output_swaps,
                "total_fidelity": self.total_fidelity,
                "total_log_fidelity": self.total_log_fidelity,
                "benchmark": {
                    "expanded_nodes": self.expanded_nodes if cfg.debug else 0,
                    "generated_nodes": self.generated_nodes if cfg.debug else 0,
                    "seconds_per_node": self.seconds_per_node,
                    "average_branching_factor": self.average_branching_factor,
                    "effective_branching_factor": self.effective_branching_factor,
                },
                "additional_gates": self.output_gates - self.input_gates,
            },
        }


def effective_branching_rate(nodes_processed: int, solution_depth: int) -> float:
    """HeuristicMapper::computeEffectiveBranchingRate (HeuristicMapper.hpp:411-435)."""
    if solution_depth == 0:
        return 0.0
    nodes_processed -= 1
    upper = float(nodes_processed) ** (1.0 / solution_depth)
    lower = upper / solution_depth
    while upper - lower > 2e-10:
        mid = (lower + upper) / 2.0
        total = 0.0
        for i in range(1, solution_depth + 1):
            total += mid ** i
        if total < nodes_processed:
            lower = mid
        else:
            upper = mid
    return (lower + upper) / 2.0


class HeuristicMapper:
    """HeuristicMapper(qc, architecture) (include/sc/heuristic/HeuristicMapper.hpp:27)."""

    def __init__(self, circuit: Circuit, architecture, lib=None, backend=None):
        self.qc = circuit
        self.architecture = architecture
        self.lib = lib if lib is not None else _abi.load()
        self._backend = backend
        self.results = MappingResults()
        self.mapped_ops: list = []
        self.initial_layout: np.ndarray | None = None  # physical -> logical (-1 free)
        self.output_permutation: np.ndarray | None = None

    # ------------------------------------------------------------------
    def _backend_or_default(self):
        if self._backend is None:
            self._backend = routing.CudaBackend(self.lib, self.architecture.handle)
        return self._backend

    def _check_parameters(self, cfg):
        """HeuristicMapper::checkParameters (:149-167), same messages."""
        if cfg.layering in (_abi.LAYERINGS["odd_gates"], _abi.LAYERINGS["qubit_triangle"]):
            name = [k for k, v in _abi.LAYERINGS.items() if v == cfg.layering][0]
            raise QmapError(f"Layering strategy {name} not suitable for heuristic mapper!")
        fid = cfg.heuristic == _abi.HEURISTICS["fidelity_best_location"]
        if fid and not self.architecture.fidelity_available:
            raise QmapError("Fidelity aware heuristic chosen, but no or insufficient "
                            "calibration data available for this architecture!")
        if fid and cfg.lookahead_heuristic != 0:
            raise QmapError("Fidelity-aware heuristics may only be used with fidelity-aware "
                            "lookahead heuristics (or no lookahead)!")

    def _static_layout(self, lc, c, t, used, n):
        """staticInitialMapping (:109-147): the 2q gates of the first layer go
        to the first free coupling edges, the other qubits follow by index."""
        loc = np.full(n, -1, dtype=np.int16)
        qub = np.full(n, -1, dtype=np.int16)
        off, idx = lc.layer_gates()
        edges = [tuple(int(v) for v in e) for e in self.architecture.edges]
        if lc.n_layers:
            for g in idx[off[0]: off[1]]:
                if c[g] < 0:
                    continue
                for q0, q1 in edges:
                    if qub[q0] == -1 and qub[q1] == -1:
                        qub[q0], qub[q1] = c[g], t[g]
                        loc[c[g]], loc[t[g]] = q0, q1
                        break
        for i in used:  # original indices with an initialLayout entry
            if i < n and loc[i] == -1:
                for j in range(n):
                    if qub[j] == -1:
                        loc[i], qub[j] = j, i
                        break
        return loc

    # ------------------------------------------------------------------
    # Configuration::dataLoggingPath (Configuration.hpp:53): a directory that
    # receives nodes_layer_<i>.csv in the DataLogger's schema (DataLogger.cpp:
    # 216-269) for the final routing pass; None = no logging
    data_logging_path = None
    node_log_capacity = 1 << 20

    def map(self, config=None, **kwargs) -> MappingResults:
        """HeuristicMapper::map (:44-107).

        Layer splits are first left to the library, which follows the
        reference bit for bit (restart on the uncleared open list).  That
        restart can return a node whose validMappedTwoQubitGates set still
        counts gates of the unsplit layer, i.e. a "valid" layout in which a
        gate of the layer is not adjacent; the reference then only asserts
        (:414) and a release build emits a CNOT on a non-edge.  Such a result
        is detected at emission and the circuit is routed again with the
        layers split on this side and a fresh open list per restart, which is
        always valid (MappingResults.rerouted_with_host_splits)."""
        cfg = config if config is not None else routing.make_config(**kwargs)
        self._check_parameters(cfg)
        try:
            return self._map_once(cfg, True)
        except InvalidPlacement:
            res = self._map_once(cfg, False)
            res.rerouted_with_host_splits = True
            return res

    def _map_once(self, cfg, library_splits: bool) -> MappingResults:
        t_start = time.perf_counter()
        n = self.architecture.n_qubits
        c, t, identity = self.qc.mapper_view(n)
        if len(self.qc.used_qubits()) > n:
            raise QmapError("More qubits in the circuit than available on the architecture.")
        used = self.qc.used_qubits()
        lc = routing.LayeredCircuit(self.lib, n, c, t, cfg.layering)
        res = MappingResults(config=cfg)
        res.input_gates = len(c)
        res.input_cnots = int((c >= 0).sum())
        res.input_single_qubit_gates = int((c < 0).sum())
        if cfg.initial_layout == _abi.LAYOUTS["identity"]:
            init = identity
        elif cfg.initial_layout == _abi.LAYOUTS["static"]:
            init = self._static_layout(lc, c, t, used, n) if lc.n_layers else None
        else:
            init = None
        start_given = init.copy() if init is not None else np.full(n, -1, dtype=np.int16)

        # ---- route.  The per-layer outputs get room for every layer a split can
        # add (a circuit never has more layers than gates), so the library
        # splits layers itself like the reference (restart on the uncleared
        # open list, splits in pseudo-routing passes) and reports the splits of
        # the final pass.  Should it still hand a split back
        # (QMG_SPLIT_LIMIT_HIT), the layer is split here and the search resumes
        # from it with the layout reached so far.
        backend = self._backend_or_default()
        extra = len(c) if cfg.automatic_layer_splits and library_splits else 0
        passes = int(cfg.iterative_bidirectional_routing_passes)
        first = 0
        lq_parts, ns_parts, swap_parts, place_parts, rec_parts = [], [], [], [], []
        logged = None  # the routed instance whose node log was written
        start_locations = None
        splits = 0
        while True:
            item = routing.RouteItem(lc, init, passes, first_layer=first)
            log_cap = self.node_log_capacity if self.data_logging_path else 0
            out = routing.route_batch(backend, n, cfg, [item], want_layer_qubits=True,
                                      lean=not (cfg.debug or log_cap), extra_layers=extra,
                                      node_log_capacity=log_cap)[0]
            if log_cap and out.status == _abi.QMG_OK and first == 0:
                from . import datalog
                datalog.write_node_logs(self.data_logging_path, out)
                logged = out
            if out.split_layers is not None and len(out.split_layers):
                lc = lc.split_copy(out.split_layers)
                splits += len(out.split_layers)
            if start_locations is None:
                start_locations = out.start_locations.copy()
            done = out.n_layers_done
            # heuristicBenchmark counts the final routing pass only (the
            # reference switches its counters off in pseudoRouteCircuit)
            res.expanded_nodes += out.expanded_nodes_final_pass
            res.generated_nodes += out.generated_nodes_final_pass
            res.expanded_nodes_all_passes += out.expanded_nodes
            res.generated_nodes_all_passes += out.generated_nodes
            lq_parts.append(out.layer_qubits[first:done].copy())
            if out.layer_n_swaps is not None:  # lean outputs: packed SWAP lists
                ns_parts.append(out.layer_n_swaps[first:done].copy())
                nsw = int(out.layer_n_swaps[first:done].sum())
                swap_parts.append(out.swaps[:nsw].copy())
            else:  # full per-layer records (debug)
                ns_parts.append(out.layers["n_swaps"][first:done].astype(np.uint16))
                swap_parts.extend(out.layer_swaps(l).copy() for l in range(first, done))
            place_parts.append(out.placements.copy())
            if cfg.debug and out.layers is not None:
                rec_parts.append(out.layers[first:done].copy())
            if out.status == _abi.QMG_OK:
                break
            if out.status != _abi.QMG_SPLIT_LIMIT_HIT:
                if out.status == _abi.QMG_NO_VIABLE_MAPPING:
                    raise QmapError("No viable mapping found.")
                raise QmapError(f"routing failed with status {out.status}")
            lc = lc.split_copy(done)  # Mapper::splitLayer
            splits += 1
            init = out.final_locations.copy()
            passes = 0
            first = done
        layer_qubits = np.concatenate(lq_parts) if lq_parts else np.zeros((0, n), np.int16)
        layer_n_swaps = np.concatenate(ns_parts) if ns_parts else np.zeros(0, np.uint16)
        swaps = np.concatenate(swap_parts) if swap_parts else np.zeros((0, 2), np.uint16)
        placements = np.concatenate(place_parts) if place_parts else np.zeros((0, 2), np.uint16)
        res.layer_splits = splits
        res.input_layers = lc.n_layers
        res.layer_n_swaps = layer_n_swaps
        # Mapper::initResults (Mapper.cpp:33-42)
        res.input_name = self.qc.name
        res.input_qubits = len(used)
        res.output_name = self.qc.name + "_mapped"
        res.output_qubits = n
        res.architecture = self.architecture.name
        if cfg.debug and rec_parts:
            res.layer_benchmark = np.concatenate(rec_parts)

        self._emit(cfg, lc, c, t, n, start_given, start_locations, layer_qubits, layer_n_swaps,
                   swaps, placements, res)
        # HeuristicMapper.cpp:432-444 (debug): totals over the final pass
        if cfg.debug and res.expanded_nodes > 0:
            res.average_branching_factor = (res.generated_nodes - lc.n_layers) / res.expanded_nodes
            if res.layer_benchmark is not None:
                for r in res.layer_benchmark:
                    e = int(r["expanded_nodes"])
                    res.effective_branching_factor += (
                        effective_branching_rate(e + 1, int(r["solution_depth"])) * (e / res.expanded_nodes))
        res.time = time.perf_counter() - t_start
        res.timeout = False
        self.results = res
        if logged is not None:
            # the rest of the DataLogger directory (architecture, layers, result)
            from . import datalog
            datalog.write_run(self.data_logging_path, self, logged, lc, c, t, res)
        return res

    # ------------------------------------------------------------------
    def _emit(self, cfg, lc, c, t, n, start_given, start_locations, layer_qubits, layer_n_swaps,
              swaps, placements, res):
        """routeCircuit (:346-519): SWAPs, gates at their locations, reversed
        CNOTs wrapped in H, counters, initial layout and output permutation.

        `initial_layout[p]` is the logical qubit that sits on physical p when
        the mapped circuit starts (-1: none).  It is obtained by tracing every
        placement back through the SWAPs emitted before it, so that replaying
        the mapped circuit from it reproduces every gate's operands exactly
        (the reference records the placement position itself, :197-208)."""
        arch = self.architecture
        directed = {tuple(int(v) for v in e) for e in arch.edges}
        ops = self.qc.ops
        off, idx = lc.layer_gates()
        out_ops = []
        to_adjust = []  # (index in out_ops, logical target) of 1q gates on unplaced qubits
        origin = np.arange(n, dtype=np.int64)  # time-0 position of what sits on p now
        initial = np.full(n, -1, dtype=np.int64)
        seen = set()

        sw_pos = 0
        for l in range(lc.n_layers):
            qubits = layer_qubits[l]
            locations = np.full(n, -1, dtype=np.int64)
            for p in range(n):
                if qubits[p] >= 0:
                    locations[qubits[p]] = p
            k = int(layer_n_swaps[l])
            layer_swaps = swaps[sw_pos: sw_pos + k]
            sw_pos += k
            emit_swaps = l != 0 or bool(cfg.swap_on_first_layer)  # (:367)
            # layout after this layer's placements, before its SWAPs
            before = qubits.astype(np.int64).copy()
            if emit_swaps:
                for a, b in layer_swaps[::-1]:
                    before[a], before[b] = before[b], before[a]
            for p in range(n):
                q = int(before[p])
                if q >= 0 and q not in seen:
                    seen.add(q)
                    initial[origin[p]] = q
            if l != 0 and cfg.add_barriers_between_layers:
                out_ops.append(Op("barrier", tuple(range(n))))
            if emit_swaps:
                for a, b in layer_swaps:
                    a, b = int(a), int(b)
                    out_ops.append(Op("swap", (a, b)))
                    origin[a], origin[b] = origin[b], origin[a]
                    res.output_swaps += 1
            for g in idx[off[l]: off[l + 1]]:
                op = ops[g]
                if c[g] < 0:
                    lt = int(t[g])
                    if locations[lt] < 0:
                        to_adjust.append((len(out_ops), lt))
                        out_ops.append(Op(op.name, (lt,), op.params))
                    else:
                        out_ops.append(Op(op.name, (int(locations[lt]),), op.params))
                else:
                    pc, pt = int(locations[c[g]]), int(locations[t[g]])
                    if (pc, pt) not in directed:
                        if (pt, pc) not in directed:
                            raise InvalidPlacement("invalid CNOT placement in the routed circuit")
                        for q in (pt, pc):
                            out_ops.append(Op("h", (q,)))
                        out_ops.append(Op(op.name, (pt, pc), op.params))
                        for q in (pc, pt):
                            out_ops.append(Op("h", (q,)))
                        res.output_direction_reverse += 1
                    else:
                        out_ops.append(Op(op.name, (pc, pt), op.params))

        final_qubits = layer_qubits[-1].astype(np.int64).copy() if lc.n_layers else \
            np.full(n, -1, dtype=np.int64)
        if not lc.n_layers:
            placed = start_locations if start_locations is not None else start_given
            for q in range(n):
                if placed[q] >= 0:
                    final_qubits[placed[q]] = q
                    initial[placed[q]] = q
        # gates on qubits that never met a 2q gate go to a qubit that is free at
        # the end; walking backwards through the SWAPs gives the position at
        # the time of each such gate (:461-508)
        if to_adjust:
            cur = final_qubits.copy()
            pending = dict()
            for pos, lt in to_adjust:
                pending.setdefault(pos, lt)
            for pos in range(len(out_ops) - 1, -1, -1):
                o = out_ops[pos]
                if o.name == "swap":
                    a, b = o.qubits
                    cur[a], cur[b] = cur[b], cur[a]
                elif pos in pending:
                    lt = pending[pos]
                    where = np.nonzero(cur == lt)[0]
                    if len(where) == 0:
                        free = int(np.nonzero(cur < 0)[0][0])
                        cur[free] = lt
                        # it ends where the free qubit ends: trace forward
                        end = free
                        for o2 in out_ops[pos + 1:]:
                            if o2.name == "swap":
                                a, b = o2.qubits
                                end = b if end == a else (a if end == b else end)
                        final_qubits[end] = lt
                        where = [free]
                    out_ops[pos] = Op(o.name, (int(where[0]),), o.params)
            for p in range(n):
                if cur[p] >= 0 and int(cur[p]) not in seen:
                    seen.add(int(cur[p]))
                    initial[p] = cur[p]
        self.mapped_ops = out_ops
        self.output_permutation = final_qubits
        self.initial_layout = initial
        self._count_gates(out_ops, directed, res)

    def _count_gates(self, out_ops, directed, res):
        """Mapper::countGates (Mapper.cpp:532-585) over the mapped circuit: a
        SWAP counts as 3 CNOTs on a bidirectional edge and as 3 CNOTs + 4 H on
        a unidirectional one; with calibration data the -log2 fidelity costs
        of all gates add up to total_log_fidelity."""
        arch = self.architecture
        fid = arch.fidelity_available
        c1 = c2 = cs = None
        if fid:
            c1, c2, cs = (arch.single_qubit_fidelity_costs(), arch.two_qubit_fidelity_costs(),
                          arch.swap_fidelity_costs())
        gates = cnots = single = 0
        log_fid = 0.0
        for o in out_ops:
            if o.name == "barrier":
                continue
            if o.name == "swap":
                a, b = o.qubits
                if fid:
                    log_fid += float(cs[a][b])
                cnots += 3
                if (a, b) in directed and (b, a) in directed:
                    gates += 3
                else:
                    gates += 7
                    single += 4
            elif len(o.qubits) == 1:
                single += 1
                gates += 1
                if fid:
                    log_fid += float(c1[o.qubits[0]])
            else:
                cnots += 1
                gates += 1
                if fid:
                    log_fid += float(c2[o.qubits[0]][o.qubits[1]])
        res.output_gates, res.output_cnots, res.output_single_qubit_gates = gates, cnots, single
        res.total_log_fidelity = log_fid
        res.total_fidelity = 2.0 ** (-log_fid)

    # ------------------------------------------------------------------
    def mapped_circuit(self) -> Circuit:
        out = Circuit(self.architecture.n_qubits, name=self.qc.name + "_mapped",
                      n_clbits=self.qc.n_clbits)
        out.ops = [o for o in self.mapped_ops if o.name != "barrier"]
        return out

    def dump_result(self, path: str):
        with open(path, "w") as f:
            f.write(self.mapped_circuit().to_qasm())


def verify_mapping(original: Circuit, mapper: HeuristicMapper) -> str:
    """Replays the mapped circuit from `mapper.initial_layout`: SWAPs permute
    the layout, every other gate must be the next unconsumed gate of the
    original circuit on the logical qubits it lands on (per-qubit order
    preserved, same name and parameters), 2q gates must sit on a coupling edge
    in the right direction (H H CX H H is a reversed CX), and the final layout
    must equal `mapper.output_permutation`.  Returns '' when the mapped circuit
    implements the original one."""
    arch = mapper.architecture
    n = arch.n_qubits
    directed = {tuple(int(v) for v in e) for e in arch.edges}
    c, t, _ = original.mapper_view(n)
    operands = [(int(c[g]), int(t[g])) if c[g] >= 0 else (int(t[g]),) for g in range(len(c))]
    queues = {}
    for g, lg in enumerate(operands):
        for q in lg:
            queues.setdefault(q, []).append(g)
    heads = {q: 0 for q in queues}
    layout = [int(v) for v in mapper.initial_layout]
    ops = [o for o in mapper.mapped_ops if o.name != "barrier"]

    def head(q):
        return queues[q][heads[q]] if q in queues and heads[q] < len(queues[q]) else None

    def consume(phys, name, params):
        lg = tuple(layout[p] for p in phys)
        if any(q < 0 for q in lg):
            return f"{name} on physical {phys}: no logical qubit there"
        g = head(lg[0])
        if g is None:
            return f"unexpected {name} on physical {phys} (logical {lg})"
        og = original.ops[g]
        if operands[g] != lg or og.name != name or tuple(og.params) != tuple(params):
            return (f"{name} on physical {phys} = logical {lg} does not match gate {g} "
                    f"({og.name}{operands[g]})")
        for x in lg:
            if head(x) != g:
                return f"gate order violated on logical qubit {x} (gate {g})"
        for x in lg:
            heads[x] += 1
        return ""

    consumed = 0
    i = 0
    while i < len(ops):
        o = ops[i]
        if o.name == "swap":
            a, b = o.qubits
            if (a, b) not in directed and (b, a) not in directed:
                return f"SWAP on ({a},{b}) is not on a coupling edge"
            layout[a], layout[b] = layout[b], layout[a]
            i += 1
            continue
        if o.name == "h":
            x = layout[o.qubits[0]]
            g = head(x) if x >= 0 else None
            plain_h = g is not None and original.ops[g].name == "h" and len(operands[g]) == 1
            wrapped = (i + 4 < len(ops) and ops[i + 1].name == "h" and len(ops[i + 2].qubits) == 2
                       and ops[i + 3].name == "h" and ops[i + 4].name == "h"
                       and {o.qubits[0], ops[i + 1].qubits[0]} == set(ops[i + 2].qubits)
                       and {ops[i + 3].qubits[0], ops[i + 4].qubits[0]} == set(ops[i + 2].qubits))
            if not plain_h and wrapped:
                pa, pb = ops[i + 2].qubits  # emitted on the edge (pa, pb): a reversed (pb, pa)
                if (pa, pb) not in directed:
                    return f"reversed CNOT on ({pa},{pb}) is not on a coupling edge"
                msg = consume((pb, pa), ops[i + 2].name, ops[i + 2].params)
                if msg:
                    return msg
                consumed += 1
                i += 5
                continue
        if len(o.qubits) == 2 and tuple(o.qubits) not in directed:
            return f"{o.name} on {o.qubits} violates the coupling direction"
        msg = consume(tuple(o.qubits), o.name, o.params)
        if msg:
            return msg
        consumed += 1
        i += 1
    if consumed != len(operands):
        return f"{len(operands) - consumed} original gates were not emitted"
    perm = mapper.output_permutation
    for p in range(n):
        if layout[p] >= 0 and perm[p] != layout[p]:
            return f"output permutation: physical {p} holds logical {layout[p]}, recorded {perm[p]}"
    return ""
