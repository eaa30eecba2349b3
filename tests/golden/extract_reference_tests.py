"""Extracts the known-answer vectors the reference's own tests hold for the
hot path into tests/golden/reference_tests.json.

Run in the build container (needs /root/reference):
    python tests/golden/extract_reference_tests.py

Sources (reference tree):
  test/sc/test_general.cpp:63-192          Dijkstra / edge-skip target matrices
  test/sc/heuristic/test_heuristic.cpp:
      276-556   OPTIMAL_SOLUTIONS: per-layer final layouts of 9 circuits mapped
                on ibmq_yorktown, ibmq_london and ibm_qx5 in that order
                (fixture :241-268: identity layout, disjoint_2q_blocks, no
                lookahead, no automatic splits; checked at :699-721 for the
                tight + principally admissible heuristics)
      776-855   generated / expanded node counts of a 3-CX circuit on a 5-ring
      869-893   generated nodes (30) of one CX(0,6) on IBM QX5
Only numbers are extracted; no reference code is copied.
"""
from __future__ import annotations

import json
import os
import re

REF = os.environ.get("QMAP_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))


def braces_after(text: str, start: int) -> tuple[str, int]:
    """the balanced {...} block starting at the first '{' at or after start"""
    i = text.index("{", start)
    depth = 0
    for j in range(i, len(text)):
        if text[j] == "{":
            depth += 1
        elif text[j] == "}":
            depth -= 1
            if depth == 0:
                return text[i:j + 1], j + 1
    raise ValueError("unbalanced braces")


def to_json(block: str):
    block = re.sub(r"//[^\n]*", "", block)
    block = block.replace("{", "[").replace("}", "]")
    block = re.sub(r",\s*\]", "]", block)
    return json.loads(block)


def value_of(body: str, name: str, names=None):
    m = re.search(r"\b" + re.escape(name) + r"\s*=\s*\{", body)
    if m is None:
        raise KeyError(name)
    block, _ = braces_after(body, m.start())
    for k, v in (names or {}).items():  # an initialiser may name an earlier value
        block = re.sub(r"\b" + re.escape(k) + r"\b", json.dumps(v).replace("[", "{").replace("]", "}"), block)
    return to_json(block)


def test_body(text: str, suite: str, name: str) -> str:
    m = re.search(r"TEST\(\s*" + suite + r"\s*,\s*" + name + r"\s*\)", text)
    block, _ = braces_after(text, m.end())
    return block


def main():
    out = {}
    gen = open(os.path.join(REF, "test/sc/test_general.cpp")).read()
    b = test_body(gen, "General", "Dijkstra")
    out["dijkstra"] = {"cm": value_of(b, "cm"), "edge_weights": value_of(b, "edgeWeights"),
                       "table": value_of(b, "targetTable1")}
    b = test_body(gen, "General", "DijkstraCNOTReversal")
    out["dijkstra_cnot_reversal"] = {"cm": value_of(b, "cm"), "edge_weights": value_of(b, "edgeWeights"),
                                     "reversal_cost": 1.0, "table": value_of(b, "targetTable2")}
    b = test_body(gen, "General", "DijkstraSkipEdges")
    table = value_of(b, "targetTable")
    out["dijkstra_skip_edges"] = {"cm": value_of(b, "cm"), "edge_weights": value_of(b, "edgeWeights"),
                                  "table": table,
                                  "edge_skip_tables": value_of(b, "edgeSkipTargetTable",
                                                               {"targetTable": table})}

    heur = open(os.path.join(REF, "test/sc/heuristic/test_heuristic.cpp")).read()
    m = re.search(r"TestHeuristics::OPTIMAL_SOLUTIONS\s*\{", heur)
    block, _ = braces_after(heur, m.end() - 1)
    entries = to_json(block)
    out["optimal_solutions"] = {
        "settings": {"initial_layout": "identity", "layering": "disjoint_2q_blocks",
                     "lookahead_heuristic": "none", "automatic_layer_splits": False},
        "architectures_in_order": ["ibmq_yorktown", "ibmq_london", "ibm_qx5"],
        "heuristics": ["gate_count_max_distance", "gate_count_sum_distance_minus_shared_swaps",
                       "gate_count_max_distance_or_sum_distance_minus_shared_swaps"],
        "layouts": {name: layouts for name, layouts in entries},
    }
    out["node_counts"] = {
        "ring5_three_cx": {
            "cm": [[0, 1], [1, 0], [1, 2], [2, 1], [2, 3], [3, 2], [3, 4], [4, 3], [4, 0], [0, 4]],
            "n": 5, "cx": [[4, 2], [3, 1], [4, 1]],
            "settings": {"heuristic": "gate_count_max_distance", "layering": "disjoint_qubits",
                         "initial_layout": "identity", "lookahead_heuristic": "none"},
            "layers": 2, "solution_depth": [1, 1], "generated_nodes": [6, 5],
            "expanded_nodes": [1, 1], "total_generated": 11, "total_expanded": 2},
        "qx5_single_cx": {
            "arch": "ibm_qx5", "n": 16, "cx": [[0, 6]],
            "settings": {"heuristic": "gate_count_max_distance", "lookahead_heuristic": "none",
                         "layering": "individual_gates", "automatic_layer_splits": False,
                         "initial_layout": "identity"},
            "generated_nodes": [30], "total_generated": 30},
    }
    path = os.path.join(HERE, "reference_tests.json")
    with open(path, "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("wrote", path, os.path.getsize(path), "bytes;",
          len(out["optimal_solutions"]["layouts"]), "circuits with optimal layouts")


if __name__ == "__main__":
    main()
