"""CPU suite: the HOST DEBUG BUILD of the routing kernel's source
(tests/emu/emu.cpp compiles mqt-qmap_b200/csrc/route_core.hpp with a 1-lane
group) against the committed outputs of the reference.  This exercises the
search logic the sm_100a kernel runs - candidate order, open-list dedup,
staged top selection, lookahead tables, fidelity scoring, placement - where
no GPU exists; the CUDA kernel itself is checked by the -m gpu tests.
Cost tables are inputs here (taken from the fixtures): the table kernels exist
only in CUDA."""
import os

import numpy as np
import pytest

import emu_backend
import parity
from qmap_b200 import _abi, routing

T = parity.golden_tables()
_backends = {}


def emu_for(name, calib=None):
    key = (name, calib)
    if key not in _backends:
        n, e = parity.read_arch_file(name)
        fid = None
        if calib:
            fid = (T[f"{name}_c1"], T[f"{name}_c2"], T[f"{name}_cswap"], T[f"{name}_fid"])
        _backends[key] = emu_backend.EmuBackend(n, e, T[f"{name}_dist"], T[f"{name}_dist_rev"],
                                                fid, pool_nodes=1 << 17)
    return _backends[key]


RUNS = [g for g in parity.golden_runs() if not g.meta["split"] and g.generated_nodes <= 50000]


@pytest.mark.parametrize("part", range(8))
def test_kernel_source_matches_reference_outputs(part):
    assert len(RUNS) > 230
    bad = []
    for g in RUNS[part::8]:
        be = emu_for(g.meta["arch"], g.meta["calibration"])
        msg, _ = parity.run_golden(be, be.lib, be.n, g)
        if msg:
            bad.append((g.label, msg))
    assert not bad, bad[:5]


def test_split_request_is_reported():
    """A search that hits the automatic split limit stops with
    QMG_SPLIT_LIMIT_HIT and names the layer (the host splits and resumes)."""
    g = next(g for g in parity.golden_runs() if g.meta["split"] and g.generated_nodes < 5000)
    be = emu_for(g.meta["arch"], g.meta["calibration"])
    msg, out = parity.run_golden(be, be.lib, be.n, g)
    assert out.status == _abi.QMG_SPLIT_LIMIT_HIT
    assert out.n_layers_done < g.n_layers


def test_lean_outputs_match_full_records():
    """layer_n_swaps (compact per-layer output) equals the n_swaps of the full
    qmg_layer_result records"""
    g = next(g for g in RUNS if g.meta["circuit"] == "qft_16.qasm" and g.meta["arch"] == "ibm_qx5")
    be = emu_for(g.meta["arch"])
    from qmap_b200 import circuit as qcirc
    kw = dict(g.meta["config"])
    cfg = routing.make_config(**kw)
    circ = qcirc.load_qasm(os.path.join(parity.CIRCUITS, g.meta["circuit"]))
    item = parity.item_for(be.lib, be.n, circ, cfg)
    full = routing.route_batch(be, be.n, cfg, [item])[0]
    lean = routing.route_batch(be, be.n, cfg, [item], lean=True)[0]
    assert lean.layers is None
    assert np.array_equal(lean.layer_n_swaps, full.layers["n_swaps"])
    assert np.array_equal(lean.swaps, full.swaps)
    for l in (0, 5, len(full.layers) - 1):
        assert np.array_equal(lean.layer_swaps(l), full.layer_swaps(l))
