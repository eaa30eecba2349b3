#!/usr/bin/env python
"""bench.py - A* expanded nodes/s of the heuristic router (BASELINE.json).

Workload at N=1 (configs[1]): examples/qft_16.qasm on IBM QX5, heuristic
gate_count_max_distance, lookahead gate_count_max_distance x15 (0.75/0.5),
individual-gates layering; a step routes a batch of B seeded relabellings of the
circuit (identity initial layout = B different initial layouts).  One step = one
launch of the routing kernel over the whole batch.

  value     expanded nodes / device time of the kernel (inputs resident in HBM)
  e2e       the same through qmg_route_batch with host buffers (staging,
            H2D, kernel, D2H inside the timed region)
  roofline  algorithmic bytes per launch / kernel time vs measured HBM peak
  cpu_baseline / --impl reference: the compiled reference (oracle/_ref) on the
            host cores over a bounded sample of the same batch
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "mqt-qmap_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

QASM = os.path.join(ROOT, "tests", "golden", "circuits", "qft_16.qasm")
ARCH = os.path.join(ROOT, "tests", "golden", "arch", "ibm_qx5.arch")
SUITE = os.path.join(ROOT, "tests", "golden", "circuits")
WORKLOAD = "qft_16 on IBM QX5 (16q), lookahead gate_count_max_distance x15, batch of seeded relabellings"
METRIC = "astar_expanded_nodes_per_sec"


def instance_perm(seed: int) -> np.ndarray:
    return np.random.default_rng(1000003 + seed).permutation(16)


def shard_seeds(rank: int, world: int, batch: int) -> range:
    """Instances of rank `rank`: every rank routes `batch` instances of its own
    (weak scaling, no data-path collective); seeds are disjoint across ranks."""
    assert 0 <= rank < world
    return range(rank * batch, (rank + 1) * batch)


def relabel_arrays(c, t, perm):
    perm = np.asarray(perm, dtype=np.int32)
    return (np.where(c >= 0, perm[np.maximum(c, 0)], -1).astype(np.int32),
            perm[t].astype(np.int32))


# ---------------------------------------------------------------------------
# CPU arm: the compiled reference on the host cores (test infrastructure)
# ---------------------------------------------------------------------------
def _ref_worker(args):
    seeds, = args
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "mqt-qmap_b200"))
    from oracle import refapi
    from qmap_b200 import circuit as qcirc
    R = refapi.Ref()
    ra = R.arch_load(ARCH)
    base = qcirc.load_qasm(QASM)
    cfg = refapi.make_config(initial_layout="identity", debug=True)
    exp = gen = swaps = 0
    t0 = time.perf_counter()
    search_s = 0.0
    c0, t0_ = base.gate_arrays()
    for s in seeds:
        c, t = relabel_arrays(c0, t0_, instance_perm(s))
        rr = R.map_gates(ra, 16, c, t, cfg)
        exp += rr.expanded_nodes
        gen += rr.generated_nodes
        swaps += rr.swaps
        search_s += rr.seconds_per_node * rr.expanded_nodes
    return exp, gen, swaps, time.perf_counter() - t0, search_s


def _ref_worker_c4(args):
    seeds, n_cx, tmpdir = args
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "mqt-qmap_b200"))
    from oracle import refapi
    from qmap_b200 import workloads
    R = refapi.Ref()
    arch_path, csv_path = workloads.c4_files(os.path.join(tmpdir, f"w{os.getpid()}"))
    ra = R.arch_load(arch_path, csv_path)  # (table build: not part of the timed search)
    cfg = refapi.make_config(debug=True, **workloads.C4_CONFIG)
    exp = gen = swaps = 0
    search_s = 0.0
    t0 = time.perf_counter()
    for s in seeds:
        c, t = workloads.c4_circuit(s, n_cx).gate_arrays()
        rr = R.map_gates(ra, 127, c, t, cfg)
        exp += rr.expanded_nodes
        gen += rr.generated_nodes
        swaps += rr.swaps
        search_s += rr.seconds_per_node * rr.expanded_nodes
    return exp, gen, swaps, time.perf_counter() - t0, search_s


def run_reference_c4(seeds, cores: int, n_cx: int):
    """Config 4 with the reference on `cores` processes; same return value as
    run_reference."""
    from concurrent.futures import ProcessPoolExecutor
    import multiprocessing as mp
    import tempfile
    chunks = [list(seeds[i::cores]) for i in range(cores)]
    chunks = [c for c in chunks if c]
    tmpdir = tempfile.mkdtemp(prefix="qmap_c4_ref_")
    t0 = time.perf_counter()
    with ProcessPoolExecutor(max_workers=len(chunks), mp_context=mp.get_context("spawn")) as ex:
        res = list(ex.map(_ref_worker_c4, [(c, n_cx, tmpdir) for c in chunks]))
    wall = time.perf_counter() - t0
    busy = max(r[3] for r in res)
    return (sum(r[0] for r in res), sum(r[1] for r in res), sum(r[2] for r in res),
            busy, wall, sum(r[4] for r in res))


def run_reference(seeds, cores: int):
    """Routes the given instances with the reference on `cores` processes.
    Returns (expanded, generated, swaps, wall seconds)."""
    from concurrent.futures import ProcessPoolExecutor
    import multiprocessing as mp
    chunks = [list(seeds[i::cores]) for i in range(cores)]
    chunks = [c for c in chunks if c]
    t0 = time.perf_counter()
    with ProcessPoolExecutor(max_workers=len(chunks), mp_context=mp.get_context("spawn")) as ex:
        res = list(ex.map(_ref_worker, [(c,) for c in chunks]))
    wall = time.perf_counter() - t0
    busy = max(r[3] for r in res)  # excludes process start-up
    return (sum(r[0] for r in res), sum(r[1] for r in res), sum(r[2] for r in res),
            busy, wall, sum(r[4] for r in res))


class WorkloadC2:
    """configs[1]: qft_16 on IBM QX5, seeded relabellings, lookahead."""
    key = "c2"
    name = WORKLOAD
    n = 16
    default_batch = 16384

    def __init__(self, args):
        self.args = args

    def cpu_sample(self, cores):
        return self.args.cpu_sample or 6000 * cores

    def ref_step_sample(self, cores):
        return self.args.cpu_sample or 600 * cores

    def run_reference(self, seeds, cores):
        return run_reference(seeds, cores)

    def architecture(self, device):
        from qmap_b200.architecture import Architecture
        return Architecture(device=device).load_coupling_map(ARCH)

    def config(self):
        from qmap_b200 import routing
        return routing.make_config(initial_layout="identity")

    def items(self, L, n, cfg, seeds):
        from qmap_b200 import routing
        from qmap_b200 import circuit as qcirc
        base = qcirc.load_qasm(QASM)
        c0, t0_ = base.gate_arrays()
        out = []
        for s in seeds:
            c, t = relabel_arrays(c0, t0_, instance_perm(s))
            lc = routing.LayeredCircuit(L, n, c, t, cfg.layering)
            out.append(routing.RouteItem(lc, np.arange(n, dtype=np.int16)))
        return out

    def options(self, opts):
        opts.searches_per_sm = self.args.searches_per_sm


class WorkloadC4:
    """configs[3]: random 127-qubit circuits on a 127-qubit heavy-hex
    architecture with (synthetic) calibration data, fidelity_best_location.
    The reference's search grows exponentially with the distance of the gate's
    qubits (10^4..10^5 nodes per layer after ~30 CX), so an instance is the
    first `--c4-cx` CX of a seeded random circuit instead of 10k CX: the
    reference routes such an instance in seconds, a 10k-CX one not at all."""
    key = "c4"
    n = 127
    default_batch = 8192

    def __init__(self, args):
        self.args = args
        self.n_cx = args.c4_cx
        self.name = (f"random 127-qubit circuits ({self.n_cx} CX + 1q gates each) on heavy-hex 127 "
                     "with synthetic calibration, heuristic fidelity_best_location, batch of seeded circuits")

    def cpu_sample(self, cores):
        return self.args.cpu_sample or 8 * cores

    def ref_step_sample(self, cores):
        return self.args.cpu_sample or 4 * cores

    def run_reference(self, seeds, cores):
        return run_reference_c4(seeds, cores, self.n_cx)

    def architecture(self, device):
        from qmap_b200 import workloads
        from qmap_b200.architecture import Architecture
        arch_path, csv_path = workloads.c4_files()
        a = Architecture(device=device).load_coupling_map(arch_path)
        a.load_properties(csv_path)
        return a

    def config(self):
        from qmap_b200 import routing, workloads
        return routing.make_config(**workloads.C4_CONFIG)

    def items(self, L, n, cfg, seeds):
        from qmap_b200 import routing, workloads
        out = []
        for s in seeds:
            circ = workloads.c4_circuit(s, self.n_cx)
            c, t, identity = circ.mapper_view(n)
            lc = routing.LayeredCircuit(L, n, c, t, cfg.layering)
            out.append(routing.RouteItem(lc, identity))
        return out

    def options(self, opts):
        # long searches: few resident searches with large node pools (the
        # budget is most of the HBM), full search variant from the start
        # (measured: 8 / 12 searches per SM -> 23.3 / 26.2 M nodes/s; 12 is what
        # the 168-register fidelity kernel keeps resident)
        opts.searches_per_sm = self.args.searches_per_sm or 12
        opts.search_variant = 1
        os.environ.setdefault("QMAP_B200_POOL_MB", "140000")


def _ref_worker_c3(args):
    seeds, = args
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "mqt-qmap_b200"))
    from oracle import refapi
    from qmap_b200 import architecture as ar
    from qmap_b200 import workloads
    R = refapi.Ref()
    n, cm = ar.heavy_hex_27()
    ra = R.arch_create(n, sorted(cm))
    cfg = refapi.make_config(debug=True, **workloads.C3_CONFIG)
    exp = gen = swaps = 0
    search_s = 0.0
    t0 = time.perf_counter()
    for s in seeds:
        _, circ = workloads.c3_instance(SUITE, s)
        c, t = circ.gate_arrays()
        rr = R.map_gates(ra, circ.n_qubits, c, t, cfg)
        exp += rr.expanded_nodes
        gen += rr.generated_nodes
        swaps += rr.swaps
        search_s += rr.seconds_per_node * rr.expanded_nodes
    return exp, gen, swaps, time.perf_counter() - t0, search_s


class WorkloadC3:
    """configs[2]: the examples/*.qasm suite (22 circuits, 3..16 qubits) on a
    27-qubit heavy-hex architecture, circuits x seeded placements, lookahead
    (the mapper's defaults)."""
    key = "c3"
    n = 27
    default_batch = 16384
    name = ("examples/*.qasm suite (22 circuits) on heavy-hex 27, default heuristic + lookahead, "
            "batch of circuits x seeded placements")

    def __init__(self, args):
        self.args = args

    def cpu_sample(self, cores):
        return self.args.cpu_sample or 1500 * cores

    def ref_step_sample(self, cores):
        return self.args.cpu_sample or 300 * cores

    def run_reference(self, seeds, cores):
        from concurrent.futures import ProcessPoolExecutor
        import multiprocessing as mp
        chunks = [c for c in (list(seeds[i::cores]) for i in range(cores)) if c]
        t0 = time.perf_counter()
        with ProcessPoolExecutor(max_workers=len(chunks), mp_context=mp.get_context("spawn")) as ex:
            res = list(ex.map(_ref_worker_c3, [(c,) for c in chunks]))
        wall = time.perf_counter() - t0
        return (sum(r[0] for r in res), sum(r[1] for r in res), sum(r[2] for r in res),
                max(r[3] for r in res), wall, sum(r[4] for r in res))

    def architecture(self, device):
        from qmap_b200 import architecture as ar
        n, cm = ar.heavy_hex_27()
        return ar.Architecture(n, cm, device=device)

    def config(self):
        from qmap_b200 import routing, workloads
        return routing.make_config(**workloads.C3_CONFIG)

    def items(self, L, n, cfg, seeds):
        from qmap_b200 import routing, workloads
        out = []
        for s in seeds:
            _, circ = workloads.c3_instance(SUITE, s)
            c, t, identity = circ.mapper_view(n)
            lc = routing.LayeredCircuit(L, n, c, t, cfg.layering)
            out.append(routing.RouteItem(lc, identity))
        return out

    def options(self, opts):
        opts.searches_per_sm = self.args.searches_per_sm


WORKLOADS = {"c2": WorkloadC2, "c3": WorkloadC3, "c4": WorkloadC4}


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# ---------------------------------------------------------------------------
class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region: NVML polled
    every few ms from a thread (nvidia-smi in a loop as a fallback)."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap"}

    def __init__(self, index: int):
        self.index = index
        self.rows = []  # (sm_mhz, max_mhz, reasons bitmask)
        self.proc = None
        self._stop = threading.Event()
        self._thread = None
        self._nvml = None

    def _poll_nvml(self):
        nv = self._nvml
        h = nv.nvmlDeviceGetHandleByIndex(self.index)
        mx = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
        while not self._stop.is_set():
            try:
                sm = float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    rs = int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
                except Exception:
                    rs = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                self.rows.append((sm, mx, rs))
            except Exception:
                break
            time.sleep(0.004)

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nvml = pynvml
            self._thread = threading.Thread(target=self._poll_nvml, daemon=True)
            self._thread.start()
            return
        except Exception:
            self._nvml = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}",
                 "--query-gpu=clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
                 "clocks_event_reasons.hw_thermal_slowdown,"
                 "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap",
                 "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self._thread = threading.Thread(target=self._read_smi, daemon=True)
            self._thread.start()
        except OSError:
            self.proc = None

    def _read_smi(self):
        bits = [0x8, 0x40, 0x20, 0x4]
        for line in self.proc.stdout:
            r = [x.strip() for x in line.split(",")]
            try:
                rs = sum(b for b, v in zip(bits, r[2:6]) if v.lower().startswith("active"))
                self.rows.append((float(r[0]), float(r[1]), rs))
            except (ValueError, IndexError):
                continue

    def stop(self):
        self._stop.set()
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        if self._thread is not None:
            self._thread.join(timeout=2)
        sm = [r[0] for r in self.rows]
        mx = max((r[1] for r in self.rows), default=0.0)
        mask = 0
        for r in self.rows:
            mask |= r[2]
        reasons = sorted(nm for bit, nm in self.REASONS.items() if mask & bit)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None,
                "reasons": reasons, "samples": len(sm)}


def algorithmic_bytes(n: int, expanded: int, generated: int) -> float:
    """per expanded node with b children, R = 4n + 64:
    R (read parent) + b*R (write children) + 32*b (queue/hash) + 16 (pop)."""
    R = 4 * n + 64
    return expanded * (R + 16) + generated * (R + 32)


def reference_arm(args, wl, cores, config):
    sample = wl.ref_step_sample(cores)
    vals = []
    last = None
    for step in range(args.warmup + args.steps):
        seeds = list(range(step * sample, (step + 1) * sample))
        exp, gen, swaps, busy, wall, search_s = wl.run_reference(seeds, cores)
        if step >= args.warmup:
            vals.append((exp, busy))
        last = (exp, gen, swaps, busy, wall, search_s)
    tot_e = sum(v[0] for v in vals)
    tot_t = sum(v[1] for v in vals)
    v = tot_e / tot_t
    desc = f"{sample} instances per step on {cores} processes (oracle/_ref = reference sources)"
    return {
        "metric": METRIC, "value": v, "unit": "nodes/s", "impl": "reference", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / max(1, args.steps),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": config,
        "cpu_baseline": {"value": v, "unit": "nodes/s", "cores": cores, "kind": "reference",
                         "sample": desc,
                         "search_loop_only_nodes_per_core_s": last[0] / last[5] if last[5] else None},
        "e2e": {"value": v, "unit": "nodes/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "circuits_per_s": sample * args.steps / tot_t,
    }


def cpu_baseline_for(wl, cores):
    sample = wl.cpu_sample(cores)
    exp, gen, swaps, busy, wall, search_s = wl.run_reference(list(range(sample)), cores)
    return {"value": exp / busy, "unit": "nodes/s", "cores": cores, "kind": "reference",
            "sample": f"{sample} seeded instances of the workload (seeds 0..{sample - 1}), "
                      f"{busy:.1f} s on {cores} processes",
            "circuits_per_s": sample / busy,
            "search_loop_only_nodes_per_core_s": exp / search_s if search_s else None}


def b200_arm(args, wl, B, steps, warmup, rank, world, local_rank, config, cpu_base):
    """One workload on this rank's GPU; returns the JSON line (rank 0) or None."""
    import torch
    import torch.distributed as dist
    from qmap_b200 import _abi, routing

    L = _abi.load()
    arch = wl.architecture(local_rank)
    h = arch.handle
    n = arch.n_qubits
    cfg = wl.config()
    seeds = shard_seeds(rank, world, B)
    items = wl.items(L, n, cfg, seeds)
    opts = _abi.QmgRouteOptions()
    wl.options(opts)
    # what routeCircuit consumes: per-layer SWAP lists + layouts (the per-call
    # benchmark records of config.debug are not requested)
    opts.lean_outputs = 1
    bufs = routing.pack(n, items, lean=True)
    batch = C.c_void_p()
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    stats = _abi.QmgRouteStats()

    def create_resident():
        # A resident batch is not re-routed by the library when a node pool
        # runs out (qmg_route_batch is): halve the resident searches, which
        # doubles the pools, until every instance fits.
        for _ in range(5):
            _abi.check(L, L.qmg_batch_create(h, C.byref(cfg), len(items), C.addressof(bufs.inst),
                                             C.byref(opts), C.byref(batch)))
            _abi.check(L, L.qmg_batch_run(batch, C.byref(stats)))
            _abi.check(L, L.qmg_batch_fetch(batch, C.addressof(bufs.res)))
            if all(int(r.status) != _abi.QMG_POOL_EXHAUSTED for r in bufs.res):
                return
            L.qmg_batch_destroy(batch)
            opts.max_searches = max(4, int(stats.grid) * 4 // 2)
        raise RuntimeError("node pools too small for the workload")

    create_resident()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        flush.fill_(1)
        torch.cuda.synchronize()
        _abi.check(L, L.qmg_batch_run(batch, C.byref(stats)))
        return stats.kernel_ms

    for _ in range(warmup):
        step()
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    kernel_ms = [step() for _ in range(steps)]
    barrier()
    clocks = sampler.stop()
    _abi.check(L, L.qmg_batch_fetch(batch, C.addressof(bufs.res)))
    outs = routing.unpack(bufs)
    bad = [o.status for o in outs if o.status != 0]
    if bad:
        raise RuntimeError(f"{len(bad)} instances failed, first status {bad[0]}")
    exp = sum(o.expanded_nodes for o in outs)
    gen = sum(o.generated_nodes for o in outs)
    swaps = sum(len(o.swaps) for o in outs)
    L.qmg_batch_destroy(batch)

    # e2e: the C-ABI call a host makes, host buffers in and out
    e2e_ms = []
    n_e2e = 2 + steps
    for i in range(n_e2e):
        b2 = routing.pack(n, items, lean=True)
        st2 = _abi.QmgRouteStats()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        _abi.check(L, L.qmg_route_batch(h, C.byref(cfg), len(items), C.addressof(b2.inst),
                                        C.addressof(b2.res), C.byref(opts), C.byref(st2)))
        dt = time.perf_counter() - t0
        if i >= 2:
            e2e_ms.append(1e3 * dt)
    h2d, d2h = int(st2.h2d_bytes), int(st2.d2h_bytes)
    e2e_launches = int(st2.kernel_launches)

    t_total = sum(kernel_ms) / 1e3
    t_e2e = sum(e2e_ms) / 1e3
    red = torch.tensor([t_total, t_e2e], dtype=torch.float64, device="cuda")
    cnt = torch.tensor([exp, gen, swaps, len(items)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(red, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    t_total, t_e2e = (float(x) for x in red.tolist())
    exp_all, gen_all, swaps_all, inst_all = (float(x) for x in cnt.tolist())
    del flush
    if rank != 0:
        return None
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f)["route_kernel" if wl.key == "c2" else "route_kernel_" + wl.key][
                "dram_bytes_per_launch"]
    except (OSError, KeyError, ValueError):
        pass
    avg_launch_s = (sum(kernel_ms) / len(kernel_ms)) / 1e3
    abytes = algorithmic_bytes(n, exp, gen)
    achieved = abytes / avg_launch_s / 1e9
    return {
        "metric": METRIC, "value": exp_all * steps / t_total, "unit": "nodes/s",
        "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": 1e3 * t_total / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
        "clocks": clocks,
        "e2e": {"value": exp_all * len(e2e_ms) / t_e2e, "unit": "nodes/s",
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": 1e3 * t_e2e / len(e2e_ms),
                "circuits_per_s": inst_all * len(e2e_ms) / t_e2e,
                "outputs": "per-layer SWAP counts + SWAP list + layouts (lean results)",
                "kernel_launches_per_call": e2e_launches,
                "last_call_ms": {"stage_h2d": float(st2.h2d_ms), "kernel": float(st2.kernel_ms),
                                 "d2h_scatter": float(st2.d2h_ms)}},
        "gpu_launches": steps,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic,
                     "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback",
                     "kernel": "route_kernel", "algorithmic_bytes_per_launch": abytes},
        "cpu_baseline": cpu_base,
        "circuits_per_s": inst_all * steps / t_total,
        "expanded_nodes_per_step": exp_all, "generated_nodes_per_step": gen_all,
        "swaps_per_step": swaps_all, "grid": int(stats.grid), "block": int(stats.block),
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS),
                    help="c2 = configs[1] (qft_16 on QX5), c3 = configs[2] (suite on heavy-hex 27), "
                         "c4 = configs[3] (127-qubit heavy-hex, fidelity-aware)")
    ap.add_argument("--batch", type=int, default=0, help="instances per GPU per step (0 = the workload's default)")
    ap.add_argument("--c4-cx", type=int, default=30, help="CX gates per config-4 instance")
    ap.add_argument("--cpu-sample", type=int, default=0, help="instances in the CPU sample (0 = auto)")
    ap.add_argument("--searches-per-sm", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true",
                    help="c2 at N=1 also measures config 4 (key 'config4'); this skips it")
    args = ap.parse_args()

    # stdout carries the JSON line and nothing else: everything libraries
    # print there while the bench runs (NCCL's version banner, ...) is sent to
    # stderr, the line itself goes to the saved descriptor
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(obj):
        sys.stdout.flush()
        os.write(real_stdout, (json.dumps(obj) + "\n").encode())

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    cores = host_cores()
    wl = WORKLOADS[args.workload](args)
    B = args.batch or wl.default_batch

    def config_of(w, b):
        return {"workload": w.name, "batch_per_gpu": b,
                "l2": "flushed between timed steps (512 MiB write)", "parallelism": f"instances x{world}"}

    if args.impl == "reference":
        if rank != 0:
            return
        emit(reference_arm(args, wl, cores, config_of(wl, B)))
        return

    # ---- B200 arm ---------------------------------------------------------
    extra = None
    if wl.key == "c2" and world == 1 and not args.no_extra:
        extra = WORKLOADS["c4"](args)
    cpu_base = cpu_extra = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # before CUDA is initialised in this process (workers are spawned)
        cpu_base = cpu_baseline_for(wl, cores)
        if extra is not None:
            try:
                cpu_extra = cpu_baseline_for(extra, cores)
            except Exception as e:  # noqa: BLE001
                cpu_extra = {"error": repr(e)}

    # one process per GPU shares the host cores for staging
    os.environ.setdefault("QMAP_B200_THREADS", str(max(2, cores // max(1, world))))
    # stdout carries the JSON line only: NCCL's banner / debug output to stderr
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    import torch
    import torch.distributed as dist

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    out = b200_arm(args, wl, B, args.steps, args.warmup, rank, world, local_rank,
                   config_of(wl, B), cpu_base)
    if extra is not None and out is not None:
        # configs[3] beside the headline: fewer steps (a step takes seconds)
        try:
            Bx = extra.default_batch
            o4 = b200_arm(args, extra, Bx, min(args.steps, 2), 3, rank, world, local_rank,
                          config_of(extra, Bx), cpu_extra)
            keep = ("value", "unit", "steps", "warmup", "ms_per_step", "config", "clocks", "e2e",
                    "gpu_launches", "roofline", "cpu_baseline", "circuits_per_s",
                    "expanded_nodes_per_step", "generated_nodes_per_step", "grid", "block")
            out["config4"] = {k: o4[k] for k in keep}
        except Exception as e:  # noqa: BLE001
            out["config4"] = {"error": repr(e)}
    if out is not None:
        emit(out)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
