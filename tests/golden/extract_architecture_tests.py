"""Extracts the closed-form vectors of the reference's Architecture tests into
tests/golden/architecture_tests.json.

Run in the build container (needs /root/reference):
    python tests/golden/extract_architecture_tests.py

Source (reference tree): test/sc/test_architecture.cpp
  274-507  FidelityDistanceBidirectionalTest      k-edge-skip tables, k = 0,1,3 (+ zero 4..6)
  509-766  FidelityDistanceSemiBidirectionalTest  the same on a partly directed tree
  768-839  FidelitySwapCostTest                   swap-cost matrix incl. > 1e20 off the edges
  841-883  FidelityDistanceCheapestPathTest       cheapest, not shortest, path
The expected values are C++ expressions in std::log2; they are evaluated here
(Python floats are the same IEEE doubles).  Only test data is extracted.
"""
from __future__ import annotations

import json
import math
import os
import re

REF = os.environ.get("QMAP_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
TESTS = ["FidelityDistanceBidirectionalTest", "FidelityDistanceSemiBidirectionalTest",
         "FidelitySwapCostTest", "FidelityDistanceCheapestPathTest"]


def body_of(text: str, name: str) -> str:
    i = text.index(f"TEST(TestArchitecture, {name})")
    j = text.find("\nTEST(", i + 1)
    return text[i: j if j > 0 else len(text)]


def strip_comments(s: str) -> str:
    s = re.sub(r"/\*.*?\*/", "", s, flags=re.S)
    return re.sub(r"//[^\n]*", "", s)


def evaluate(expr: str):
    expr = " ".join(expr.split()).replace("std::log2", "log2").replace("{", "[").replace("}", "]")
    return eval(expr, {"__builtins__": {}}, {"log2": math.log2})  # noqa: S307 (test data)


def main():
    text = open(os.path.join(REF, "test", "sc", "test_architecture.cpp")).read()
    out = []
    for name in TESTS:
        b = strip_comments(body_of(text, name))
        cm = evaluate(re.search(r"CouplingMap\s+cm\s*=\s*(\{.*?\});", b, re.S).group(1))
        n = max(max(e) for e in cm) + 1
        q1 = {int(m.group(1)): float(m.group(2))
              for m in re.finditer(r'setSingleQubitErrorRate\((\d+),\s*"x",\s*([0-9.]+)\)', b)}
        cx = [[int(m.group(1)), int(m.group(2)), float(m.group(3))]
              for m in re.finditer(r"setTwoQubitErrorRate\((\d+),\s*(\d+),\s*([0-9.]+)\)", b)]
        mats = {m.group(1): evaluate(m.group(2))
                for m in re.finditer(r"const Matrix\s+(\w+)\s*=\s*(\{.*?\});", b, re.S)}
        tables = []
        for m in re.finditer(r"matrixNear\(\s*architecture\.getFidelityDistanceTable\((\d*)\),\s*(\w+),\s*([0-9.e-]+)\)", b):
            tables.append({"skip": int(m.group(1) or 0), "table": mats[m.group(2)], "tol": float(m.group(3))})
        near = []
        for m in re.finditer(r"EXPECT_NEAR\((\w+)\[(\d+)\]\[(\d+)\],\s*(.*?),\s*(tolerance|[0-9.e-]+)\);", b, re.S):
            near.append({"what": m.group(1), "i": int(m.group(2)), "j": int(m.group(3)),
                         "value": evaluate(m.group(4)),
                         "tol": 1e-6 if m.group(5) == "tolerance" else float(m.group(5))})
        huge = [{"what": m.group(1), "i": int(m.group(2)), "j": int(m.group(3))}
                for m in re.finditer(r"EXPECT_GT\((\w+)\[(\d+)\]\[(\d+)\],\s*1e20\)", b)]
        out.append({"test": name, "n": n, "coupling_map": cm,
                    "single_qubit_error": [q1[q] for q in range(n)], "two_qubit_error": cx,
                    "fidelity_distance_tables": tables, "near": near, "greater_than_1e20": huge})
        print(name, "n", n, "tables", [t["skip"] for t in tables], "near", len(near), "huge", len(huge))
    with open(os.path.join(HERE, "architecture_tests.json"), "w") as f:
        json.dump({"source": "test/sc/test_architecture.cpp:274-883 (cda-tum/mqt-qmap)", "tests": out}, f, indent=1)


if __name__ == "__main__":
    main()
