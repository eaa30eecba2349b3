#!/usr/bin/env python
"""profiles/traffic.json from the round's ncu captures of route_kernel: what
bench.py reads for roofline.traffic and roofline_l2 (measured DRAM / L2 bytes
per expanded node, L2 peak, issue-slot and instruction-cache figures).

  python tools/make_traffic.py <c2.ncu-rep> <expanded c2> <c4.ncu-rep> <expanded c4> <tag>
"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def metrics(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h, u, v = rows[0], rows[1], rows[2]
    m = {}
    for i, x in enumerate(h):
        try:
            m[x] = (float(v[i]), u[i])
        except ValueError:
            pass
    return m


def tobytes(val, unit):
    return val * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]


def entry(rep, exp, workload, tag):
    m = metrics(rep)
    dr, dw = tobytes(*m["dram__bytes_read.sum"]), tobytes(*m["dram__bytes_write.sum"])
    l2 = m["lts__t_sectors.sum"][0] * 32
    t = m["gpu__time_duration.sum"]
    t = t[0] * {"ms": 1e-3, "s": 1, "us": 1e-6}[t[1]]
    slices = round(m["lts__cycles_elapsed.sum"][0] / m["lts__cycles_elapsed.avg"][0])
    clk = m["lts__cycles_elapsed.avg.per_second"][0]
    g = lambda k: m[k][0] if k in m else None  # noqa: E731
    return {
        "workload": workload, "expanded_nodes": exp, "kernel_s_under_ncu": t,
        "dram_bytes_per_launch": dr + dw, "dram_bytes_per_expanded_node": (dr + dw) / exp,
        "l2_bytes_per_launch": l2, "l2_bytes_per_expanded_node": l2 / exp,
        "l2_peak_gbs": slices * 32 * clk,
        "l2_peak_note": f"{slices} L2 slices x 32 B/cycle x {clk:.3f} GHz (slice data path); ncu "
                        f"lts__t_sectors pct_of_peak = {g('lts__t_sectors.avg.pct_of_peak_sustained_elapsed'):.2f} %",
        "l2_hit_pct": g("lts__t_sector_hit_rate.pct"),
        "issue_slot_busy_pct": g("smsp__issue_active.avg.pct_of_peak_sustained_active"),
        "warps_active_per_scheduler": g("smsp__warps_active.avg.per_cycle_active"),
        "lanes_per_instruction": g("smsp__thread_inst_executed_per_inst_executed.ratio"),
        "warp_instructions": g("smsp__inst_executed.sum"),
        "warp_instructions_per_expanded_node": g("smsp__inst_executed.sum") / exp,
        "sm_icache_hit_pct": g("sm__icc_request_hit_rate.pct"),
        "gpc_icache_requests": g("gcc__cache_requests_type_instruction.sum"),
        "gpc_icache_requests_pct_of_peak": g("gcc__cache_requests_type_instruction.sum.pct_of_peak_sustained_elapsed"),
        "gpc_icache_requests_per_expanded_node": g("gcc__cache_requests_type_instruction.sum") / exp,
        "local_load_sectors": g("l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum"),
        "local_store_sectors": g("l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum"),
        "local_l1_miss_bytes": 32 * (g("l1tex__t_sectors_pipe_lsu_mem_local_op_ld_lookup_miss.sum") +
                                     g("l1tex__t_sectors_pipe_lsu_mem_local_op_st_lookup_miss.sum")),
        "registers_per_thread": g("launch__registers_per_thread"),
        "source": f"ncu --set full --clock-control none, one launch ({tag}; summaries: profiles/r2_*)",
    }


def main():
    c2, e2, c4, e4, tag = sys.argv[1], int(sys.argv[2]), sys.argv[3], int(sys.argv[4]), sys.argv[5]
    res = {
        "route_kernel_c2": entry(c2, e2, "config 2: 16384 relabelled qft_16 instances on IBM QX5 "
                                         "(bench.py --workload c2)", tag),
        "route_kernel_c4": entry(c4, e4, "config 4: 2048 instances x 30 CX on heavy-hex 127 (bench.py --workload c4 "
                                         "--batch 2048); the per-node figures are applied to the full step", tag),
    }
    with open(os.path.join(ROOT, "profiles", "traffic.json"), "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
