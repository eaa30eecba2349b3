"""-m "not gpu": bench.py's reference arm runs on the host alone (the
reference's sources compiled under oracle/_ref), so its JSON line is checked
here; the B200 arm must fail loudly without a GPU (no CPU fallback)."""
import json
import os
import subprocess
import sys

import pytest

import parity

BENCH = os.path.join(parity.ROOT, "bench.py")
needs_ref = pytest.mark.skipif(
    not os.path.exists(os.path.join(parity.ROOT, "oracle", "_ref", "libqmapref.so")),
    reason="oracle/_ref not built")


def _run(*args, timeout=600):
    return subprocess.run([sys.executable, BENCH, *args], capture_output=True, text=True,
                          timeout=timeout, cwd=parity.ROOT)


@needs_ref
@pytest.mark.parametrize("workload,sample", [("c2", "64"), ("c4", "8")])
def test_reference_arm_prints_the_contract_line(workload, sample):
    r = _run("--impl", "reference", "--workload", workload, "--steps", "1", "--warmup", "0",
             "--cpu-sample", sample)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference"
    assert d["metric"] == "astar_expanded_nodes_per_sec" and d["unit"] == "nodes/s"
    assert d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["n_gpus"] == 1 and d["steps"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0
    assert d["config"]["workload"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and cb["sample"]
    assert cb["value"] == d["value"]
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0
    assert e["unit"] == d["unit"] and e["value"] > 0
    # nothing of the product library on this arm
    assert "libqmap_b200" not in r.stderr


def test_b200_arm_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = _run("--workload", "c2", "--steps", "1", "--warmup", "1", "--no-cpu-baseline",
             "--no-extra", "--batch", "8")
    assert r.returncode != 0
    assert not [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
