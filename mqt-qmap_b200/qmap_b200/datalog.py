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
