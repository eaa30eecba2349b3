"""Search-node logs in the reference's DataLogger format
(src/sc/DataLogger.cpp:216-269): one file `nodes_layer_<i>.csv` per layer, one
line per generated node

    nodeId;parentId;costFixed;costHeur;lookaheadPenalty;validMapping;depth;q0,q1,...;a b,c d,...

The device log (routing.RouteOutput.node_log) holds, per node, the costs, ids,
depth and the SWAP that made it; layouts and SWAP sequences are rebuilt here
from the ancestors and the layer's root layout.  Runs in which a layer was
split are not supported (node ids restart inside one layer's file)."""
from __future__ import annotations

import os

import numpy as np


def _num(x: float) -> str:
    # std::ostream << double with the default precision (6 significant digits)
    return f"{float(x):g}"


def layer_lines(out, n_layers=None) -> list:
    """[[csv line, ...] per layer] of the instance's node log."""
    if out.node_log is None or out.layer_root_qubits is None:
        raise ValueError("the instance was routed without a node log")
    if out.n_node_log > len(out.node_log):
        raise ValueError(f"node log cut: {out.n_node_log} nodes generated, room for {len(out.node_log)}")
    if out.split_layers is not None and len(out.split_layers):
        raise ValueError("node logs of runs with layer splits are not supported")
    nl = n_layers if n_layers is not None else len(out.layer_root_qubits)
    lines = [[] for _ in range(nl)]
    nodes = {}
    cur_layer = -1
    for r in out.node_log:
        layer = int(r["layer"])
        if layer != cur_layer:
            nodes = {}
            cur_layer = layer
        if r["root"]:
            q = [int(x) for x in out.layer_root_qubits[layer]]
            swaps = []
        else:
            pq, ps = nodes[int(r["parent"])]
            a, b = int(r["swap_first"]), int(r["swap_second"])
            q = list(pq)
            q[a], q[b] = q[b], q[a]
            swaps = ps + [(a, b)]
        nodes[int(r["id"])] = (q, swaps)
        lines[layer].append(";".join([
            str(int(r["id"])), str(int(r["parent"])), _num(r["cost_fixed"]), _num(r["cost_heur"]),
            _num(r["lookahead_penalty"]), str(int(r["valid"])), str(int(r["depth"])),
            ",".join(str(x) for x in q), ",".join(f"{a} {b}" for a, b in swaps)]))
    return lines


def write_node_logs(directory: str, out, n_layers=None):
    """nodes_layer_<i>.csv for every layer of the instance."""
    os.makedirs(directory, exist_ok=True)
    for i, ls in enumerate(layer_lines(out, n_layers)):
        with open(os.path.join(directory, f"nodes_layer_{i}.csv"), "w") as f:
            f.write("".join(line + "\n" for line in ls))


def parse_nodes_csv(path: str):
    """The reference's nodes_layer_<i>.csv as a list of dicts (as
    test/sc/heuristic/test_heuristic.cpp:170-222 parses it)."""
    rows = []
    with open(path) as f:
        for line in f:
            line = line.rstrip("\n")
            if not line:
                continue
            p = line.split(";")
            rows.append(dict(id=int(p[0]), parent=int(p[1]), cost_fixed=float(p[2]), cost_heur=float(p[3]),
                             lookahead_penalty=float(p[4]), valid=int(p[5]), depth=int(p[6]),
                             qubits=[int(x) for x in p[7].split(",")] if p[7] else [],
                             swaps=[tuple(int(y) for y in x.split()) for x in p[8].split(",")] if len(p) > 8 and p[8] else []))
    return rows


# ---------------------------------------------------------------------------
# the other files of the DataLogger directory (DataLogger.cpp:52-186, :270-296),
# which the reference's SearchVisualizer reads next to the node files
# ---------------------------------------------------------------------------
def _qasm3_ops(ops) -> str:
    """One statement per gate in the style of the reference's OpenQASM 3 dump
    (mqt-core; `cx q[0], q[1];`).  The exact text of mqt-core is not pinned."""
    lines = []
    for o in ops:
        p = ""
        if o.params:
            p = "(" + ",".join(repr(float(x)) for x in o.params) + ")"
        lines.append(f"{o.name}{p} " + ", ".join(f"q[{q}]" for q in o.qubits) + ";\n")
    return "".join(lines)


def architecture_json(arch) -> dict:
    """DataLogger::logArchitecture (:52-80)."""
    n = arch.n_qubits
    j = {"name": arch.name, "nqubits": n,
         "coupling_map": [[int(a), int(b)] for a, b in arch.coupling_map],
         "distances": [[float(x) for x in row] for row in arch.distance_table(True)]}
    if arch.fidelity_available:
        p = arch.properties
        single = [1.0] * n
        for q in sorted(p.single_qubit_error):
            single[q] = 1.0 - p.average_single_qubit_error_rate(q)
        two = [[0.0] * n for _ in range(n)]
        for a, b in arch.coupling_map:
            if (a, b) in p.two_qubit_error:
                two[a][b] = 1.0 - p.two_qubit_error[(a, b)]
        j["fidelity"] = {
            "single_qubit_fidelities": single,
            "two_qubit_fidelities": two,
            "single_qubit_fidelity_costs": [float(x) for x in arch.single_qubit_fidelity_costs()],
            "two_qubit_fidelity_costs": [[float(x) for x in r] for r in arch.two_qubit_fidelity_costs()],
            "swap_fidelity_costs": [[float(x) for x in r] for r in arch.swap_fidelity_costs()],
            "fidelity_distances": [[[float(x) for x in r] for r in t] for t in arch.fidelity_distance_tables()],
        }
    return j


def layer_json(n, layer_ops, g2, g1, root_qubits, record, final_qubits, final_swaps) -> dict:
    """DataLogger::logFinalizeLayer (:98-186): g2 = [(q1, q2, forward, backward)],
    g1 = [(qubit, multiplicity)] of the layer, `record` its qmg_layer_result."""
    single = [0] * n
    for q, m in g1:
        single[int(q)] = int(m)
    return {
        "qasm": _qasm3_ops(layer_ops),
        "two_qubit_multiplicity": [{"q1": int(a), "q2": int(b), "forward": int(f), "backward": int(r)}
                                   for a, b, f, r in g2],
        "single_qubit_multiplicity": single,
        "initial_layout": [int(x) for x in root_qubits],
        "final_node_id": int(record["result_node_id"]),
        "final_cost_fixed": float(record["cost_fixed"]),
        "final_cost_heur": float(record["cost_heur"]),
        "final_lookahead_penalty": float(record["lookahead_penalty"]),
        "final_layout": [int(x) for x in final_qubits],
        "final_swaps": [[int(a), int(b)] for a, b in final_swaps],
        "final_search_depth": int(record["solution_depth"]),
    }


def layer_benchmark_json(record, effective_branching_factor: float) -> dict:
    """LayerHeuristicBenchmarkInfo::json (MappingResults.hpp:75-92)."""
    e, g = int(record["expanded_nodes"]), int(record["generated_nodes"])
    return {
        "expanded_nodes": e,
        "generated_nodes": g,
        "expanded_nodes_after_first_solution": int(record["expanded_nodes_after_first_solution"]),
        "expanded_nodes_after_optimal_solution": int(record["expanded_nodes_after_optimal_solution"]),
        "solution_nodes": int(record["solution_nodes"]),
        "solution_nodes_after_optimal_solution": int(record["solution_nodes_after_optimal_solution"]),
        "solution_depth": int(record["solution_depth"]),
        "seconds_per_node": 0.0,
        "average_branching_factor": (g - 1) / e if e > 0 else 0.0,
        "effective_branching_factor": effective_branching_factor,
        "early_termination": bool(record["early_termination"]),
    }


def _layer_multiplicities(lay):
    """[(g2, g1) per layer] of a qmg_layers view: g2 = [(q1, q2, forward,
    backward)], g1 = [(qubit, multiplicity)]."""
    import ctypes as C
    nl = int(lay.n_layers)
    o2 = np.ctypeslib.as_array(C.cast(lay.gate2q_offsets, C.POINTER(C.c_uint32)), (nl + 1,))
    o1 = np.ctypeslib.as_array(C.cast(lay.gate1q_offsets, C.POINTER(C.c_uint32)), (nl + 1,))
    g2 = (np.ctypeslib.as_array(C.cast(lay.gates2q, C.POINTER(C.c_uint16)), (int(o2[-1]) * 4,)).reshape(-1, 4)
          if nl and o2[-1] else np.zeros((0, 4), np.uint16))
    g1 = (np.ctypeslib.as_array(C.cast(lay.gates1q, C.POINTER(C.c_uint16)), (int(o1[-1]) * 2,)).reshape(-1, 2)
          if nl and o1[-1] else np.zeros((0, 2), np.uint16))
    return [(g2[o2[l]:o2[l + 1]], g1[o1[l]:o1[l + 1]]) for l in range(nl)]


def write_run(directory: str, mapper, out, lc, control, target, results):
    """Everything DataLogger leaves in its directory besides the node files:
    architecture.json, input.qasm, layer_<i>.json, output.qasm and
    mapping_result.json (DataLogger.cpp:52-186, :270-296; HeuristicMapper.cpp:
    45-48, :102-106).  `out` is the routed instance with per-layer records."""
    import json
    from .circuit import Op
    from .mapper import effective_branching_rate
    os.makedirs(directory, exist_ok=True)
    arch = mapper.architecture
    n = arch.n_qubits

    def dump(name, obj):
        with open(os.path.join(directory, name), "w") as f:
            json.dump(obj, f, indent=2)

    dump("architecture.json", architecture_json(arch))
    ops = mapper.qc.ops
    view = [Op(ops[g].name, ((int(control[g]), int(target[g])) if control[g] >= 0 else (int(target[g]),)),
               ops[g].params) for g in range(len(ops))]
    head = 'OPENQASM 3.0;\ninclude "stdgates.inc";\n'
    with open(os.path.join(directory, "input.qasm"), "w") as f:
        f.write(head + f"qubit[{results.input_qubits}] q;\n" + _qasm3_ops(view))
    off, idx = lc.layer_gates()
    mult = _layer_multiplicities(lc.layers)
    layers_json = []
    for l in range(lc.n_layers):
        rec = out.layers[l]
        g2, g1 = mult[l]
        dump(f"layer_{l}.json",
             layer_json(n, [view[g] for g in idx[off[l]: off[l + 1]]], g2, g1, out.layer_root_qubits[l], rec,
                        out.layer_qubits[l], out.layer_swaps(l)))
        e = int(rec["expanded_nodes"])
        layers_json.append(layer_benchmark_json(rec, effective_branching_rate(e + 1, int(rec["solution_depth"]))))
    with open(os.path.join(directory, "output.qasm"), "w") as f:
        f.write(head + f"qubit[{n}] q;\n" + _qasm3_ops([o for o in mapper.mapped_ops if o.name != "barrier"]))
    j = results.json()
    if results.config.debug:
        j["statistics"]["benchmark"]["layers"] = layers_json
    dump("mapping_result.json", j)
