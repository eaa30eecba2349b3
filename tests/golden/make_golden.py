"""Generates the routing / table golden vectors from the REFERENCE ITSELF
(oracle/_ref/libqmapref.so = the reference's own src/sc sources compiled by
oracle/Makefile; needs /root/reference, so run it in the build container):

    python tests/golden/make_golden.py

Outputs (committed):
  tests/golden/route_golden.json   manifest: one entry per case (architecture,
                                   circuit, configuration keywords, totals)
  tests/golden/route_golden.npz    per case i: c{i}_calls (qmg_layer_result per
                                   final-pass aStarMap call), c{i}_qubits
                                   (layout after each call), c{i}_swaps
  tests/golden/tables_golden.npz   distance tables of every .arch fixture and
                                   the fidelity cost / edge-skip tables of
                                   ibmq_london with its calibration CSV
The test-suite checks the CPU restatement (oracle/qmap_oracle.cpp), the host
debug build of the kernel source and the CUDA library against these files.
"""
from __future__ import annotations

import glob
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "mqt-qmap_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

from oracle import refapi  # noqa: E402
from qmap_b200 import circuit as qcirc  # noqa: E402

CIRCUITS = os.path.join(HERE, "circuits")
ARCHS = os.path.join(HERE, "arch")

HEUR = ["gate_count_max_distance", "gate_count_sum_distance",
        "gate_count_sum_distance_minus_shared_swaps",
        "gate_count_max_distance_or_sum_distance_minus_shared_swaps"]
LAYERINGS = ["individual_gates", "disjoint_qubits", "disjoint_2q_blocks"]
LAYOUTS = ["dynamic", "identity"]
LOOK = ["none", "gate_count_max_distance", "gate_count_sum_distance"]
ET = ["expanded_nodes", "expanded_nodes_after_first_solution",
      "expanded_nodes_after_current_optimal_solution", "solution_nodes",
      "solution_nodes_after_current_optimal_solution"]


def case_list():
    files = sorted(glob.glob(os.path.join(CIRCUITS, "*.qasm")), key=os.path.getsize)
    nq = {f: qcirc.load_qasm(f).n_qubits for f in files}
    cases = []
    k = 0
    # every circuit on every gate-count architecture, rotating the settings
    for arch, n in (("ibm_qx4", 5), ("ibmq_yorktown", 5), ("ibm_qx5", 16), ("ibmq_tokyo", 20),
                    ("linear_10qubit", 10)):
        for f in files:
            if nq[f] > n:
                continue
            for rep in range(2):
                kw = dict(heuristic=HEUR[k % 4], layering=LAYERINGS[(k // 4) % 3],
                          initial_layout=LAYOUTS[(k // 2) % 2], lookahead_heuristic=LOOK[(k // 3) % 3])
                if k % 7 == 3:
                    kw["iterative_bidirectional_routing_passes"] = 1
                if k % 11 == 5:
                    kw["nr_lookaheads"] = 4
                    kw["first_lookahead_factor"] = 0.9
                    kw["lookahead_factor"] = 0.7
                cases.append((arch, None, os.path.basename(f), kw))
                k += 1
    # the bench configurations
    cases.append(("ibm_qx4", None, "4gt13_92.qasm", dict(lookahead_heuristic="none")))
    cases.append(("ibm_qx5", None, "qft_16.qasm", dict()))
    # early termination (non-tight search needs the fidelity heuristic; with the
    # tight ones the limit only matters for the goal found at the top)
    for i, et in enumerate(ET):
        for f in ("rd32-v0_66.qasm", "4mod5-v0_20.qasm"):
            cases.append(("ibmq_london", "ibmq_london.csv", f,
                          dict(heuristic="fidelity_best_location", lookahead_heuristic="none",
                               layering=LAYERINGS[i % 3], early_termination=et,
                               early_termination_limit=2 + i)))
    # fidelity-aware routing
    for f in files:
        if nq[f] > 5:
            continue
        for lay in LAYERINGS:
            for il in LAYOUTS:
                cases.append(("ibmq_london", "ibmq_london.csv", os.path.basename(f),
                              dict(heuristic="fidelity_best_location", lookahead_heuristic="none",
                                   layering=lay, initial_layout=il)))
    # automatic layer splits (restart on the stale open list)
    for arch, n in (("ibm_qx5", 16), ("ibmq_tokyo", 20)):
        for f in files:
            if nq[f] > n or nq[f] < 5:
                continue
            cases.append((arch, None, os.path.basename(f),
                          dict(layering="disjoint_qubits", automatic_layer_splits_node_limit=3)))
            cases.append((arch, None, os.path.basename(f),
                          dict(layering="disjoint_2q_blocks", automatic_layer_splits_node_limit=5,
                               heuristic="gate_count_sum_distance",
                               iterative_bidirectional_routing_passes=1)))
    return cases


def main():
    R = refapi.Ref()
    archs = {}
    manifest, blobs = [], {}
    for i, (arch, calib, qasm, kw) in enumerate(case_list()):
        key = (arch, calib)
        if key not in archs:
            archs[key] = R.arch_load(os.path.join(ARCHS, arch + ".arch"),
                                     os.path.join(ARCHS, calib) if calib else None)
        ra = archs[key]
        cfg = refapi.make_config(debug=True, **kw)
        rr = R.map_qasm(ra, os.path.join(CIRCUITS, qasm), cfg)
        nl = rr.n_layers
        keep = np.nonzero((rr.call_flags & 2) == 0)[0][-nl:] if nl else np.zeros(0, dtype=np.int64)
        calls = rr.calls[keep].copy()
        swaps = []
        off = 0
        for c in calls:
            s = rr.swap_pairs[c["swap_offset"]: c["swap_offset"] + c["n_swaps"]]
            swaps.append(s)
            c["swap_offset"] = off
            off += len(s)
        blobs[f"c{i}_calls"] = calls
        blobs[f"c{i}_qubits"] = rr.qubits_after[keep]
        blobs[f"c{i}_swaps"] = (np.concatenate(swaps) if swaps else np.zeros((0, 2))).astype(np.uint16)
        manifest.append({"arch": arch, "calibration": calib, "circuit": qasm, "config": kw,
                         "layers": int(nl), "split": bool((rr.call_flags & 2).any()),
                         "expanded_nodes": int(rr.expanded_nodes),
                         "generated_nodes": int(rr.generated_nodes), "swaps": int(rr.swaps)})
    np.savez_compressed(os.path.join(HERE, "route_golden.npz"), **blobs)
    with open(os.path.join(HERE, "route_golden.json"), "w") as f:
        json.dump(manifest, f, indent=0)

    tables = {}
    for af in sorted(glob.glob(os.path.join(ARCHS, "*.arch"))):
        name = os.path.basename(af)[:-5]
        ra = R.arch_load(af)
        tables[f"{name}_dist"] = ra.distance_table(False)
        tables[f"{name}_dist_rev"] = ra.distance_table(True)
    ra = R.arch_load(os.path.join(ARCHS, "ibmq_london.arch"), os.path.join(ARCHS, "ibmq_london.csv"))
    tables["ibmq_london_c1"] = ra.single_qubit_fidelity_costs()
    tables["ibmq_london_c2"] = ra.two_qubit_fidelity_costs()
    tables["ibmq_london_cswap"] = ra.swap_fidelity_costs()
    tables["ibmq_london_fid"] = ra.fidelity_distance_tables()
    np.savez_compressed(os.path.join(HERE, "tables_golden.npz"), **tables)
    nsplit = sum(m["split"] for m in manifest)
    print(len(manifest), "cases,", nsplit, "with layer splits,",
          os.path.getsize(os.path.join(HERE, "route_golden.npz")), "bytes")


if __name__ == "__main__":
    main()
