"""-m "not gpu": the compiled C++ drop-in (integration/GpuHeuristicMapper.hpp
around the reference's own map()) on the HOST build of the kernel source
(oracle/_ref/libqmapdropin_emu.so, `make -C oracle dropin_emu`): the same
comparison tests/test_gpu_dropin.py makes through libqmap_b200.so on the GPU,
on every third golden configuration, plus Configuration::dataLoggingPath
through the drop-in against the files plain HeuristicMapper writes."""
import filecmp
import json
import os
import subprocess
import tempfile

import pytest

import parity
from oracle import refapi
from test_gpu_dropin import _same

DROPIN_EMU = os.path.join(parity.ROOT, "oracle", "_ref", "libqmapdropin_emu.so")
pytestmark = pytest.mark.skipif(not os.path.isdir("/root/reference/src/sc") and not os.path.exists(DROPIN_EMU),
                                reason="needs the reference's sources (or the prebuilt library)")


@pytest.fixture(scope="module")
def dropin():
    if os.path.isdir("/root/reference/src/sc"):
        subprocess.check_call(["make", "-s", "-C", os.path.join(parity.ROOT, "oracle"), "ref", "dropin_emu"])
    return refapi.Ref(DROPIN_EMU)


def _archs(R, dropin, cache, m):
    key = (m["arch"], m["calibration"])
    if key not in cache:
        files = (os.path.join(parity.ARCHS, m["arch"] + ".arch"),
                 os.path.join(parity.ARCHS, m["calibration"]) if m["calibration"] else None)
        cache[key] = (R.arch_load(*files), dropin.arch_load(*files))
    return cache[key]


def test_golden_runs_through_the_drop_in_on_the_host_build(dropin):
    R = parity.ref()
    cache = {}
    runs = parity.golden_runs()[::3]  # (all 314 pass too; 5 minutes on the host build)
    for g in runs:
        ra, da = _archs(R, dropin, cache, g.meta)
        cfg = refapi.make_config(debug=True, **g.meta["config"])
        path = os.path.join(parity.CIRCUITS, g.meta["circuit"])
        _same(dropin.map_qasm(da, path, cfg), R.map_qasm(ra, path, cfg), g.label)
    assert len(runs) > 100


@pytest.mark.parametrize("arch,calib,heuristic", [
    ("ibm_qx5", None, "gate_count_max_distance"),
    ("ibm_qx5", None, "gate_count_sum_distance_minus_shared_swaps"),
    ("ibmq_london", "ibmq_london.csv", "fidelity_best_location")])
def test_data_logging_through_the_drop_in(dropin, arch, calib, heuristic):
    """DataLogger files (DataLogger.cpp:101-269) written by the reference's
    map() through the drop-in equal those of plain HeuristicMapper: every
    nodes_layer_<i>.csv and layer_<i>.json; mapping_result.json up to the
    timing fields."""
    R = parity.ref()
    files = (os.path.join(parity.ARCHS, arch + ".arch"),
             os.path.join(parity.ARCHS, calib) if calib else None)
    ra, da = R.arch_load(*files), dropin.arch_load(*files)
    n_files = 0
    for circuit in parity.DATALOG_CIRCUITS[:4]:
        if arch == "ibmq_london" and circuit not in ("3_17_13", "ex-1_166", "ham3_102"):
            continue
        path = os.path.join(parity.CIRCUITS, circuit + ".qasm")
        cfg = refapi.make_config(**dict(parity.DATALOG_CONFIG, heuristic=heuristic))
        d_ref = tempfile.mkdtemp(prefix="qmap_log_ref_") + os.sep
        d_dev = tempfile.mkdtemp(prefix="qmap_log_dev_") + os.sep
        a = dropin.map_qasm(da, path, cfg, d_dev)
        b = R.map_qasm(ra, path, cfg, d_ref)
        _same(a, b, circuit)
        names = sorted(os.listdir(d_ref))
        assert sorted(os.listdir(d_dev)) == names
        for f in names:
            if f.startswith("nodes_layer_") or f.startswith("layer_") or f in ("architecture.json", "input.qasm",
                                                                               "output.qasm"):
                assert filecmp.cmp(d_ref + f, d_dev + f, shallow=False), (circuit, f)
                n_files += 1
        ja, jb = json.load(open(d_dev + "mapping_result.json")), json.load(open(d_ref + "mapping_result.json"))
        for j in (ja, jb):  # (timing fields differ)
            j["statistics"].pop("mapping_time", None)
            j["statistics"].get("benchmark", {}).pop("seconds_per_node", None)
            for lb in j["statistics"].get("benchmark", {}).get("layers", []):
                lb.pop("seconds_per_node", None)
        assert ja == jb, circuit
    assert n_files > 10
