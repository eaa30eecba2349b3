#!/usr/bin/env python
"""Key metrics of the first kernel in an .ncu-rep (ncu --set full).
  python tools/ncu_metrics.py gpurun_out/x.ncu-rep"""
import csv
import io
import subprocess
import sys

WANT = ['gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size',
        'launch__shared_mem_per_block_dynamic', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem',
        'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio',
        'smsp__warps_active.avg.per_cycle_active', 'smsp__warps_eligible.avg.per_cycle_active',
        'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_bytes.sum', 'lts__t_sectors.sum', 'lts__t_sector_hit_rate.pct',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_sectors_srcunit_tex_op_read.sum', 'lts__t_sectors_srcunit_tex_op_write.sum',
        'l1tex__t_sector_hit_rate.pct', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum', 'l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum',
        'l1tex__t_sectors_pipe_lsu_mem_local_op_ld_lookup_miss.sum',
        'l1tex__t_sectors_pipe_lsu_mem_local_op_st_lookup_miss.sum',
        'smsp__sass_inst_executed_op_local_ld.sum', 'smsp__sass_inst_executed_op_local_st.sum',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__cycles_elapsed.max',
        'sm__icc_requests_lookup_hit.sum', 'sm__icc_requests_lookup_miss.sum',
        'gcc__cache_requests_type_instruction.sum',
        'gcc__cache_requests_type_instruction.sum.pct_of_peak_sustained_elapsed',
        'sm__icc_request_hit_rate.pct', 'sm__icc_requests.sum',
        'lts__t_sectors.avg.pct_of_peak_sustained_elapsed']


def main():
    out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    for i, h in enumerate(hdr):
        if h in WANT or (len(sys.argv) > 2 and sys.argv[2] in h):
            print(f"{h},{units[i]},{vals[i]}")
    # stall reasons per issued instruction
    for i, h in enumerate(hdr):
        if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio"):
            print(f"{h},{units[i]},{vals[i]}")


if __name__ == "__main__":
    main()
