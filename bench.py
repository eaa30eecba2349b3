#!/usr/bin/env python
"""bench.py - A* expanded nodes/s of the heuristic router (BASELINE.json).

Workload at N=1 (configs[3], the one the target is stated on): seeded random
127-qubit circuits on a 127-qubit heavy-hex architecture with synthetic
calibration data, heuristic fidelity_best_location.  configs[1] (qft_16 on IBM
QX5, lookahead) and configs[2] (the examples suite on heavy-hex 27) are
measured beside it at N=1 (keys "config2", "config3") or alone with
--workload c2 / c3.  One step = one launch of the routing kernel over the
whole batch of instances.

  value     expanded nodes / device time of the kernel (inputs resident in HBM)
  e2e       the same through qmg_route_batch with host buffers: staging, H2D,
            kernel, D2H of SWAP lists and layouts inside the timed region
            (layers already built: "route only")
  e2e_full  the same starting from the callers' gate lists: createLayers for
            the whole batch (qmg_circuit_batch_create) is inside the timed
            region as well
  roofline  algorithmic bytes per launch / kernel time vs measured HBM peak;
            roofline_l2: measured L2 bytes per launch (ncu) / kernel time
  cpu_baseline / --impl reference: the compiled reference (oracle/_ref) on the
            host cores over a bounded sample of the same instances; value =
            expanded nodes / time in the search loop (the reference's own
            definition, HeuristicMapper.cpp:549,603,643,659-660) x cores
  parity    the instances the CPU sample shares with the batch: expanded,
            generated and SWAP counts equal instance by instance
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


def mapper_view(c, t, n_physical: int):
    """What the reference's mapper makes of a gate list before routing
    (qmap_b200.circuit.Circuit.mapper_view, vectorised): idle qubits are
    stripped - operand q becomes its rank among the used qubits (Mapper.cpp:
    44-53, 166-172) - and InitialLayout::Identity places logical i on physical
    i for every original used index i (HeuristicMapper.cpp:176-183)."""
    present = np.zeros(int(max(t.max(), c.max())) + 1 if len(t) else 1, dtype=bool)
    present[t] = True
    present[c[c >= 0]] = True
    used = np.nonzero(present)[0].astype(np.int32)
    lut = np.full(int(used.max()) + 1 if len(used) else 1, -1, dtype=np.int32)
    lut[used] = np.arange(len(used), dtype=np.int32)
    t2 = lut[t].astype(np.int32)
    c2 = np.where(c >= 0, lut[np.maximum(c, 0)], -1).astype(np.int32)
    identity = np.full(n_physical, -1, dtype=np.int16)
    u = used[used < n_physical]
    identity[u] = u
    return c2, t2, identity


def log(msg: str):
    """progress on stderr (stdout carries the JSON line only)"""
    print(f"[bench {time.strftime('%H:%M:%S')}] {msg}", file=sys.stderr, flush=True)


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# ---------------------------------------------------------------------------
# CPU arm: the compiled reference on the host cores (test infrastructure).
# One pool of worker processes; every worker loads the architecture once and
# routes instances it takes from a shared queue (dynamic hand-out, so a few
# long instances do not leave the other workers idle).
# ---------------------------------------------------------------------------
_W = {}


def _ref_init(key, opts):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "mqt-qmap_b200"))
    from oracle import refapi
    wl = WORKLOADS[key](argparse.Namespace(**opts))
    R = refapi.Ref()
    _W["wl"], _W["R"], _W["arch"] = wl, R, wl.ref_architecture(R)
    _W["cfg"] = refapi.make_config(debug=True, **wl.config_kwargs())


def _ref_ready(_):
    return os.getpid()


def _ref_route(seed):
    wl, R = _W["wl"], _W["R"]
    nq, c, t, passes = wl.instance_gates(seed)
    cfg = _W["cfg"]
    cfg.iterative_bidirectional_routing_passes = passes
    rr = R.map_gates(_W["arch"], nq, c, t, cfg)
    # (expanded, generated: the final routing pass, as heuristicBenchmark
    # counts them; search seconds = seconds_per_node x expanded)
    return (seed, rr.expanded_nodes, rr.generated_nodes, rr.swaps,
            rr.seconds_per_node * rr.expanded_nodes, rr.seconds)


def run_reference(wl, seeds, cores: int):
    """Routes the given instances with the reference on `cores` processes.
    Returns a dict of per-instance arrays (in `seeds` order) and times."""
    import multiprocessing as mp
    seeds = list(seeds)
    log(f"reference: {len(seeds)} instances of {wl.key} on {cores} processes")
    t0 = time.perf_counter()
    with mp.get_context("spawn").Pool(cores, _ref_init, (wl.key, vars(wl.args))) as pool:
        pool.map(_ref_ready, range(cores * 4), chunksize=1)  # workers initialised
        t1 = time.perf_counter()
        res = list(pool.imap_unordered(_ref_route, seeds, chunksize=max(1, len(seeds) // (cores * 16))))
        wall = time.perf_counter() - t1
    log(f"reference: done in {wall:.1f} s (+ {t1 - t0:.1f} s start-up)")
    by = {r[0]: r for r in res}
    arr = lambda k, dt: np.array([by[s][k] for s in seeds], dtype=dt)  # noqa: E731
    return {"seeds": seeds, "expanded": arr(1, np.int64), "generated": arr(2, np.int64),
            "swaps": arr(3, np.int64), "search_s": float(arr(4, np.float64).sum()),
            "map_s": float(arr(5, np.float64).sum()), "wall_s": wall,
            "startup_s": t1 - t0}


# ---------------------------------------------------------------------------
# workloads: every instance is (circuit gate list, identity start layout,
# pseudo-routing passes), a pure function of its seed, so that the B200 arm,
# the CPU baseline and the parity tests route exactly the same instances
# ---------------------------------------------------------------------------
class Workload:
    key = ""
    n = 0
    default_batch = 0

    def __init__(self, args):
        self.args = args

    def config_kwargs(self) -> dict:
        raise NotImplementedError

    def instance_gates(self, seed):
        """(number of circuit qubits, control, target, bidirectional passes)"""
        raise NotImplementedError

    def config(self):
        from qmap_b200 import routing
        return routing.make_config(**self.config_kwargs())

    def options(self, opts):
        opts.searches_per_sm = self.args.searches_per_sm


class WorkloadC2(Workload):
    """configs[1]: qft_16 on IBM QX5, lookahead; instances = seeded relabellings
    of the circuit routed from the identity layout (= different initial
    layouts of the same circuit)."""
    key = "c2"
    n = 16
    default_batch = 16384
    name = ("qft_16 on IBM QX5 (16q), lookahead gate_count_max_distance x15, batch of seeded "
            "relabellings routed from the identity layout")
    _base = None

    def config_kwargs(self):
        return dict(initial_layout="identity")

    def cpu_sample(self, cores):
        return self.args.cpu_sample or 6000 * cores

    def ref_step_sample(self, cores):
        return self.args.cpu_sample or 600 * cores

    def instance_gates(self, seed):
        if WorkloadC2._base is None:
            from qmap_b200 import circuit as qcirc
            WorkloadC2._base = qcirc.load_qasm(QASM).gate_arrays()
        c, t = relabel_arrays(*WorkloadC2._base, instance_perm(seed))
        return 16, c, t, 0

    def ref_architecture(self, R):
        return R.arch_load(ARCH)

    def architecture(self, device):
        from qmap_b200.architecture import Architecture
        return Architecture(device=device).load_coupling_map(ARCH)


class WorkloadC4(Workload):
    """configs[3]: random 127-qubit circuits on a 127-qubit heavy-hex
    architecture with (synthetic) calibration data, fidelity_best_location.
    The reference's search grows exponentially with the distance of the gate's
    qubits (10^4..10^5 nodes per layer after ~30 CX, minutes per instance after
    ~60), so an instance is the first `--c4-cx` CX of a seeded random 10k-CX
    circuit: the reference routes such an instance in seconds, the full
    circuit not at all."""
    key = "c4"
    n = 127
    default_batch = 32768

    def __init__(self, args):
        super().__init__(args)
        self.n_cx = args.c4_cx
        self.name = (f"first {self.n_cx} CX (+ 1q gates) of seeded random 10k-CX 127-qubit circuits on "
                     "heavy-hex 127 with synthetic calibration (seed 127), heuristic "
                     "fidelity_best_location, lookahead none, initial layout identity "
                     "(random circuits = random initial layouts), batch of seeded circuits")

    def config_kwargs(self):
        from qmap_b200 import workloads
        return dict(workloads.C4_CONFIG)

    def cpu_sample(self, cores):
        return self.args.cpu_sample or 8 * cores

    def ref_step_sample(self, cores):
        return self.args.cpu_sample or 4 * cores

    def instance_gates(self, seed):
        from qmap_b200 import workloads
        c, t = workloads.c4_circuit(seed, self.n_cx).gate_arrays()
        return 127, c, t, 0

    def ref_architecture(self, R):
        import tempfile
        from qmap_b200 import workloads
        arch_path, csv_path = workloads.c4_files(tempfile.mkdtemp(prefix="qmap_c4_ref_"))
        return R.arch_load(arch_path, csv_path)  # (table build: outside the timed search)

    def architecture(self, device):
        from qmap_b200 import workloads
        from qmap_b200.architecture import Architecture
        arch_path, csv_path = workloads.c4_files()
        a = Architecture(device=device).load_coupling_map(arch_path)
        a.load_properties(csv_path)
        return a

    def options(self, opts):
        # long searches: few resident searches with large node pools (the
        # budget is most of the HBM), full search variant from the start
        opts.searches_per_sm = self.args.searches_per_sm or 12
        opts.search_variant = 1
        os.environ.setdefault("QMAP_B200_POOL_MB", "140000")


class WorkloadC3(Workload):
    """configs[2]: the examples/*.qasm suite on a 27-qubit heavy-hex
    architecture: circuits x seeded placements x {forward only, + 1
    bidirectional pass}, the mapper's defaults (lookahead)."""
    key = "c3"
    n = 27
    default_batch = 16384

    def __init__(self, args):
        super().__init__(args)
        from qmap_b200 import workloads
        self.dir = workloads.c3_directory(ROOT)
        k = len(workloads.c3_arrays(self.dir))
        self.full = k > 100
        if self.full:
            # BASELINE.md section 3.3 at its stated size: every circuit x 64
            # seeded placements x {forward only, + 1 bidirectional pass}
            self.default_batch = k * 64 * 2
        self.name = (f"examples/*.qasm suite ({k} circuits, {os.path.relpath(self.dir, ROOT)}) on heavy-hex 27, "
                     "default heuristic + lookahead, circuits x seeded placements x "
                     "{forward, +1 bidirectional pass}")

    def config_kwargs(self):
        from qmap_b200 import workloads
        return dict(workloads.C3_CONFIG)

    def cpu_sample(self, cores):
        # (the full suite holds circuits of 10^5 gates: whole rounds over it)
        return self.args.cpu_sample or (151 * 8 if self.full else 1000 * cores)

    def ref_step_sample(self, cores):
        return self.args.cpu_sample or (151 * 4 if self.full else 200 * cores)

    def instance_gates(self, seed):
        from qmap_b200 import workloads
        return workloads.c3_instance_arrays(self.dir, seed)

    def ref_architecture(self, R):
        from qmap_b200 import architecture as ar
        n, cm = ar.heavy_hex_27()
        return R.arch_create(n, sorted(cm))

    def architecture(self, device):
        from qmap_b200 import architecture as ar
        n, cm = ar.heavy_hex_27()
        return ar.Architecture(n, cm, device=device)


WORKLOADS = {"c2": WorkloadC2, "c3": WorkloadC3, "c4": WorkloadC4}


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


def _ref_summary(r, cores, sample_desc):
    """cpu_baseline object from run_reference's result.  value: the
    reference's own node rate (expanded nodes / seconds inside the search
    loop, summed over the instances) times the cores that ran in parallel."""
    exp = int(r["expanded"].sum())
    n = len(r["seeds"])
    per_core = exp / r["search_s"] if r["search_s"] > 0 else None
    return {
        "value": per_core * cores if per_core else None, "unit": "nodes/s", "cores": cores,
        "kind": "reference", "sample": sample_desc,
        "definition": "sum(expandedNodes) / sum(search-loop seconds) x cores "
                      "(HeuristicMapper.cpp:549,603,643,659-660)",
        "search_loop_nodes_per_core_s": per_core,
        "whole_map_nodes_per_s": exp / r["map_s"] * cores if r["map_s"] > 0 else None,
        "circuits_per_s": n / r["map_s"] * cores if r["map_s"] > 0 else None,
        "wall_nodes_per_s": exp / r["wall_s"], "wall_circuits_per_s": n / r["wall_s"],
        "wall_s": r["wall_s"], "expanded": exp, "generated": int(r["generated"].sum()),
        "swaps": int(r["swaps"].sum()), "instances": n,
    }


def reference_arm(args, wl, cores, config):
    """--impl reference: the reference's own sources (oracle/_ref) on all host
    cores; each step routes a bounded sample of the workload's instances
    (seeds step*S .. (step+1)*S-1, the first S of which are the first S
    instances of the B200 arm's batch)."""
    sample = wl.ref_step_sample(cores)
    tot = None
    wall = 0.0
    for step in range(args.warmup + args.steps):
        seeds = list(range(step * sample, (step + 1) * sample))
        r = run_reference(wl, seeds, cores)
        if step >= args.warmup:
            wall += r["wall_s"]
            if tot is None:
                tot = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in r.items()}
            else:
                for k in ("expanded", "generated", "swaps"):
                    tot[k] = np.concatenate([tot[k], r[k]])
                tot["seeds"] = tot["seeds"] + r["seeds"]
                for k in ("search_s", "map_s", "wall_s"):
                    tot[k] += r[k]
    desc = (f"{sample} instances per step (seeds step*{sample}..; the workload's own instances) on "
            f"{cores} processes, oracle/_ref = the reference's sources compiled")
    cb = _ref_summary(tot, cores, desc)
    v = cb["value"]
    return {
        "metric": METRIC, "value": v, "unit": "nodes/s", "impl": "reference", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * wall / max(1, args.steps),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": config, "cpu_baseline": cb,
        "e2e": {"value": cb["whole_map_nodes_per_s"], "unit": "nodes/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0,
                "definition": "expanded nodes / the reference's own map() time (results.time) x cores"},
        "circuits_per_s": cb["circuits_per_s"],
    }


def cpu_baseline_for(wl, cores):
    sample = wl.cpu_sample(cores)
    r = run_reference(wl, list(range(sample)), cores)
    cb = _ref_summary(r, cores, f"{sample} seeded instances of the workload (seeds 0..{sample - 1} = the "
                                f"first {sample} instances of the batch), {r['wall_s']:.1f} s on {cores} processes")
    return cb, r


def build_instances(L, wl, n, cfg, seeds):
    """The instances `seeds` of workload `wl` as the mapper sees them: the
    concatenated gate lists (offsets, control, target), their layers
    (routing.CircuitBatch) and one routing.RouteItem each."""
    from qmap_b200 import routing
    gl = [wl.instance_gates(s) for s in seeds]
    views = [mapper_view(g[1], g[2], n) for g in gl]
    sizes = [len(v[0]) for v in views]
    off = np.concatenate(([0], np.cumsum(sizes))).astype(np.uint64)
    ctrl = np.ascontiguousarray(np.concatenate([v[0] for v in views]), dtype=np.int32)
    tgt = np.ascontiguousarray(np.concatenate([v[1] for v in views]), dtype=np.int32)
    cb = routing.CircuitBatch(L, n, off, ctrl, tgt, cfg.layering)
    items = [routing.RouteItem(cb[i], views[i][2], int(gl[i][3])) for i in range(len(seeds))]
    return cb, items, off, ctrl, tgt


def emitted_swaps(out, cfg) -> int:
    """results.output.swaps of the reference: the SWAPs emitted, i.e. not those
    of the first layer unless swapOnFirstLayer (HeuristicMapper.cpp:369-384)."""
    first = 0 if cfg.swap_on_first_layer or len(out.layer_n_swaps) == 0 else int(out.layer_n_swaps[0])
    return len(out.swaps) - first


# ---------------------------------------------------------------------------
# cost tables of config 5's architecture (32 x 32 grid, n = 1024): the table
# kernels against Architecture::createDistanceTable of the reference, which
# runs in a host process of its own while the GPU arm works
# ---------------------------------------------------------------------------
def _tables_ref_worker(path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "mqt-qmap_b200"))
    from oracle import refapi
    from qmap_b200 import architecture as ar
    n, cm = ar.grid(32, 32)
    R = refapi.Ref()
    t0 = time.perf_counter()
    ra = R.arch_create(n, sorted(cm))
    dt = time.perf_counter() - t0
    np.savez(path, dist=ra.distance_table(False), dist_rev=ra.distance_table(True), seconds=dt)


def tables_start():
    import multiprocessing as mp
    import tempfile
    path = os.path.join(tempfile.mkdtemp(prefix="qmap_tables_"), "ref.npz")
    p = mp.get_context("spawn").Process(target=_tables_ref_worker, args=(path,))
    p.start()
    return p, path


def tables_finish(handle, device):
    from qmap_b200 import architecture as ar
    p, path = handle
    n, cm = ar.grid(32, 32)
    t0 = time.perf_counter()
    a = ar.Architecture(n, cm, device=device)
    info = a.info()
    wall = time.perf_counter() - t0
    d, dr = a.distance_table(False), a.distance_table(True)
    log("tables: waiting for the reference's createDistanceTable (n = 1024)")
    p.join(timeout=600)
    out = {"architecture": "32 x 32 grid (n = 1024, 3968 coupling entries)",
           "what": "Architecture::createDistanceTable: Dijkstra::buildTable + buildSingleEdgeSkipTable",
           "gpu_kernel_ms": float(info.table_build_ms), "gpu_call_ms": 1e3 * wall,
           "cpu_ms": None, "parity": None}
    if p.exitcode == 0 and os.path.exists(path):
        r = np.load(path)
        out["cpu_ms"] = 1e3 * float(r["seconds"])
        out["cpu"] = "the reference's sources (oracle/_ref), 1 core, beside the GPU arm"
        out["parity"] = bool(np.array_equal(r["dist"], d) and np.array_equal(r["dist_rev"], dr))
        out["speedup_kernel"] = out["cpu_ms"] / out["gpu_kernel_ms"]
    else:
        p.kill()
    return out


def _traffic(key):
    """Measured per-launch figures of the routing kernel from the committed ncu
    capture (profiles/traffic.json): DRAM bytes, L2 bytes, L2 peak."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get("route_kernel_" + key, {})
    except (OSError, ValueError):
        return {}


def b200_arm(args, wl, B, steps, warmup, rank, world, local_rank, config, cpu):
    """One workload on this rank's GPU; returns the JSON line (rank 0) or None.
    cpu = (cpu_baseline object, run_reference result) or None."""
    import torch
    import torch.distributed as dist
    from qmap_b200 import _abi, routing

    L = _abi.load()
    log(f"{wl.key}: architecture + tables")
    arch = wl.architecture(local_rank)
    h = arch.handle
    n = arch.n_qubits
    cfg = wl.config()
    seeds = list(shard_seeds(rank, world, B))
    # ---- the callers' gate lists (host arrays): not part of any timed region
    log(f"{wl.key}: {len(seeds)} gate lists")
    cb, items, off, ctrl, tgt = build_instances(L, wl, n, cfg, seeds)
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

    log(f"{wl.key}: resident batch")
    create_resident()
    log(f"{wl.key}: first run {stats.kernel_ms:.1f} ms; timing {warmup}+{steps} steps")

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
    ex = np.sort(np.array([o.expanded_nodes for o in outs]))[::-1]
    log(f"{wl.key}: expanded nodes per instance: max {ex[:5].tolist()}, mean {ex.mean():.0f}, "
        f"median {int(np.median(ex))}, sum {int(ex.sum())}; per resident search {ex.sum() / (stats.grid * 4):.0f}")
    exp = sum(o.expanded_nodes for o in outs)
    gen = sum(o.generated_nodes for o in outs)
    swaps = sum(len(o.swaps) for o in outs)
    L.qmg_batch_destroy(batch)

    log(f"{wl.key}: kernel steps {[round(x, 1) for x in kernel_ms]} ms; verify + e2e")
    # ---- every routed circuit replayed (qmg_circuit_batch_verify, host)
    t0 = time.perf_counter()
    verdict = cb.verify(arch.edges, bufs.res, len(items), fidelity_aware=cfg.heuristic == 4)
    verify_s = time.perf_counter() - t0
    parity = {"verified_instances": int((verdict == 1).sum()), "invalid": int((verdict == 0).sum()),
              "not_checkable": int((verdict == 2).sum()), "verify_s": verify_s,
              "verifier": "qmg_circuit_batch_verify: SWAPs on edges, every 2q gate on adjacent qubits "
                          "after its layer's SWAPs, final layout reached"}
    # ---- the instances shared with the CPU sample, one by one
    if cpu is not None:
        r = cpu[1]
        idx = {s: i for i, s in enumerate(seeds)}
        sh = [s for s in r["seeds"] if s in idx]
        ge = np.array([outs[idx[s]].expanded_nodes_final_pass for s in sh])
        gg = np.array([outs[idx[s]].generated_nodes_final_pass for s in sh])
        gs = np.array([emitted_swaps(outs[idx[s]], cfg) for s in sh])
        pos = [r["seeds"].index(s) for s in sh]
        re_, rg, rs = r["expanded"][pos], r["generated"][pos], r["swaps"][pos]
        parity.update({
            "shared_instances": len(sh),
            "expanded_equal": bool(np.array_equal(ge, re_)),
            "generated_equal": bool(np.array_equal(gg, rg)),
            "swaps_equal": bool(np.array_equal(gs, rs)),
            "sum_expanded": [int(ge.sum()), int(re_.sum())],
            "sum_generated": [int(gg.sum()), int(rg.sum())],
            "sum_swaps": [int(gs.sum()), int(rs.sum())],
            "compared": "per instance, B200 vs reference: expandedNodes / generatedNodes of the final "
                        "routing pass, SWAPs emitted (results.output.swaps)"})
    parity["ok"] = bool(parity["invalid"] == 0 and parity["not_checkable"] == 0 and
                        all(parity.get(k, True) for k in ("expanded_equal", "generated_equal", "swaps_equal")))

    # ---- e2e: the C-ABI call a host makes, host buffers in and out (layers
    # built beforehand: "route only")
    def timed_calls(with_layers):
        ms = []
        last = None
        for i in range(2 + steps):
            b2 = routing.pack(n, items, lean=True)
            st2 = _abi.QmgRouteStats()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            if with_layers:
                # createLayers of the whole batch from the gate lists
                cb2 = routing.CircuitBatch(L, n, off, ctrl, tgt, cfg.layering)
                cb2.fill_instances(b2.inst, len(items))
            _abi.check(L, L.qmg_route_batch(h, C.byref(cfg), len(items), C.addressof(b2.inst),
                                            C.addressof(b2.res), C.byref(opts), C.byref(st2)))
            dt = time.perf_counter() - t0
            if i >= 2:
                ms.append(1e3 * dt)
            last = st2
            if with_layers:
                bad2 = sum(1 for r in b2.res if int(r.status) != 0)
                if bad2:
                    raise RuntimeError(f"e2e_full: {bad2} instances failed")
                del cb2
        return ms, last

    e2e_ms, st2 = timed_calls(False)
    full_ms, st3 = timed_calls(True)
    h2d, d2h = int(st2.h2d_bytes), int(st2.d2h_bytes)

    t_total = sum(kernel_ms) / 1e3
    t_e2e = sum(e2e_ms) / 1e3
    t_full = sum(full_ms) / 1e3
    red = torch.tensor([t_total, t_e2e, t_full], dtype=torch.float64, device="cuda")
    cnt = torch.tensor([exp, gen, swaps, len(items)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(red, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    t_total, t_e2e, t_full = (float(x) for x in red.tolist())
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
    tr = _traffic(wl.key)
    avg_launch_s = (sum(kernel_ms) / len(kernel_ms)) / 1e3
    abytes = algorithmic_bytes(n, exp, gen)
    achieved = abytes / avg_launch_s / 1e9
    out = {
        "metric": METRIC, "value": exp_all * steps / t_total, "unit": "nodes/s",
        "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": 1e3 * t_total / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
        "clocks": clocks,
        "e2e": {"value": exp_all * len(e2e_ms) / t_e2e, "unit": "nodes/s",
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": 1e3 * t_e2e / len(e2e_ms),
                "circuits_per_s": inst_all * len(e2e_ms) / t_e2e,
                "scope": "route only: qmg_route_batch on host buffers (layers built before); staging, H2D, "
                         "kernel, D2H of SWAP lists + layouts inside the timed region",
                "kernel_launches_per_call": int(st2.kernel_launches),
                "layers_in_large_pools": int(st2.big_pool_layers),
                "retried_instances": int(st2.retried_instances),
                "last_call_ms": {"stage_h2d": float(st2.h2d_ms), "kernel": float(st2.kernel_ms),
                                 "d2h_scatter": float(st2.d2h_ms)}},
        "e2e_full": {"value": exp_all * len(full_ms) / t_full, "unit": "nodes/s",
                     "ms_per_step": 1e3 * t_full / len(full_ms),
                     "circuits_per_s": inst_all * len(full_ms) / t_full,
                     "scope": "from gate lists: createLayers of the whole batch (qmg_circuit_batch_create, "
                              "host threads) + qmg_route_batch incl. SWAP-list fetch inside the timed region",
                     "h2d_bytes_per_step": int(st3.h2d_bytes), "d2h_bytes_per_step": int(st3.d2h_bytes),
                     "host_threads": int(os.environ.get("QMAP_B200_THREADS", "0")) or None},
        "gpu_launches": steps,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak,
                     # measured DRAM bytes per expanded node (ncu) x this step's nodes
                     "traffic": (tr["dram_bytes_per_expanded_node"] * exp
                                 if tr.get("dram_bytes_per_expanded_node") else None),
                     "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback",
                     "kernel": "route_kernel", "algorithmic_bytes_per_launch": abytes,
                     "traffic_source": tr.get("source")},
        "parity": parity,
        "cpu_baseline": cpu[0] if cpu is not None else None,
        "circuits_per_s": inst_all * steps / t_total,
        "expanded_nodes_per_step": exp_all, "generated_nodes_per_step": gen_all,
        "swaps_per_step": swaps_all, "grid": int(stats.grid), "block": int(stats.block),
    }
    if tr.get("l2_bytes_per_expanded_node") and tr.get("l2_peak_gbs"):
        # L2 bytes moved per expanded node are a property of the kernel on this
        # workload (ncu capture); the rate is this run's
        l2 = tr["l2_bytes_per_expanded_node"] * exp / avg_launch_s / 1e9
        out["roofline_l2"] = {"bound": "l2", "achieved": l2, "peak": tr["l2_peak_gbs"], "unit": "GB/s",
                              "frac": l2 / tr["l2_peak_gbs"],
                              "l2_bytes_per_expanded_node": tr["l2_bytes_per_expanded_node"],
                              "l2_hit_pct": tr.get("l2_hit_pct"),
                              "issue_slot_busy_pct": tr.get("issue_slot_busy_pct"),
                              "source": tr.get("source")}
    if tr.get("gpc_icache_requests_per_expanded_node") and tr.get("gpc_icache_requests_pct_of_peak"):
        # what actually bounds the kernel: the request
        # rate of the GPC-level instruction cache.  Requests per expanded node
        # and the peak rate (requests/s at 100 %) come from the ncu capture;
        # the achieved rate is this run's.
        peak_rps = tr["gpc_icache_requests"] / tr["kernel_s_under_ncu"] / (tr["gpc_icache_requests_pct_of_peak"] / 100)
        ach = tr["gpc_icache_requests_per_expanded_node"] * exp / avg_launch_s
        out["roofline_ifetch"] = {"bound": "gpc instruction cache request rate", "achieved": ach / 1e9,
                                  "peak": peak_rps / 1e9, "unit": "G requests/s", "frac": ach / peak_rps,
                                  "requests_per_expanded_node": tr["gpc_icache_requests_per_expanded_node"],
                                  "sm_icache_hit_pct": tr.get("sm_icache_hit_pct"),
                                  "warp_instructions_per_expanded_node": tr.get("warp_instructions_per_expanded_node"),
                                  "source": tr.get("source")}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS),
                    help="c4 = configs[3] (127-qubit heavy-hex, fidelity-aware; the headline), "
                         "c2 = configs[1] (qft_16 on QX5), c3 = configs[2] (suite on heavy-hex 27)")
    ap.add_argument("--batch", type=int, default=0, help="instances per GPU per step (0 = the workload's default)")
    ap.add_argument("--c4-cx", type=int, default=30, help="CX gates per config-4 instance")
    ap.add_argument("--cpu-sample", type=int, default=0, help="instances in the CPU sample (0 = auto)")
    ap.add_argument("--searches-per-sm", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true",
                    help="at N=1 the headline workload is followed by the other two configs "
                         "(keys 'config2', 'config3'); this skips them")
    args = ap.parse_args()
    import faulthandler
    faulthandler.dump_traceback_later(240, repeat=True, file=sys.stderr)

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
                "instances": f"seeds rank*{b} .. rank*{b}+{b - 1} of the workload",
                "l2": "flushed between timed steps (512 MiB write)", "parallelism": f"instances x{world}"}

    if args.impl == "reference":
        if rank != 0:
            return
        emit(reference_arm(args, wl, cores, config_of(wl, B)))
        return

    # ---- B200 arm ---------------------------------------------------------
    extras = []
    if world == 1 and not args.no_extra:
        extras = [WORKLOADS[k](args) for k in ("c2", "c3", "c4") if k != wl.key]
    cpu = {}
    tables = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline and not args.no_extra:
        tables = tables_start()
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # before CUDA is initialised in this process (workers are spawned)
        for w in [wl] + extras:
            try:
                cpu[w.key] = cpu_baseline_for(w, cores)
            except Exception as e:  # noqa: BLE001
                cpu[w.key] = None
                print(f"cpu baseline of {w.key} failed: {e!r}", file=sys.stderr)

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
                   config_of(wl, B), cpu.get(wl.key))
    ok = out is None or out["parity"]["ok"]
    for w in extras:
        if out is None:
            break
        # the other configs beside the headline (same keys, fewer steps)
        name = {"c2": "config2", "c3": "config3", "c4": "config4"}[w.key]
        try:
            Bx = w.default_batch
            o = b200_arm(args, w, Bx, min(args.steps, 2), 3, rank, world, local_rank,
                         config_of(w, Bx), cpu.get(w.key))
            keep = ("value", "unit", "steps", "warmup", "ms_per_step", "config", "clocks", "e2e", "e2e_full",
                    "gpu_launches", "roofline", "roofline_l2", "roofline_ifetch", "parity", "cpu_baseline", "circuits_per_s",
                    "expanded_nodes_per_step", "generated_nodes_per_step", "grid", "block")
            out[name] = {k: o[k] for k in keep if k in o}
            ok = ok and o["parity"]["ok"]
        except Exception as e:  # noqa: BLE001
            out[name] = {"error": repr(e)}
    if out is not None and tables is not None:
        try:
            out["tables"] = tables_finish(tables, local_rank)
            ok = ok and out["tables"]["parity"] is not False
        except Exception as e:  # noqa: BLE001
            out["tables"] = {"error": repr(e)}
    if out is not None:
        emit(out)
    if world > 1:
        dist.destroy_process_group()
    if not ok:
        print("PARITY FAILED: see the 'parity' keys of the line", file=sys.stderr)
        sys.exit(3)


if __name__ == "__main__":
    main()
