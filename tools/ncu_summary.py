#!/usr/bin/env python
"""Key counters of every kernel in an ncu report (ncu -i <rep> --page raw --csv), one block per launch.
usage: tools/ncu_summary.py <report.ncu-rep> [cells_per_launch]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
cells = float(sys.argv[2]) if len(sys.argv) > 2 else None
out = open(rep).read() if rep.endswith(".csv") else subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout   # (a raw page exported on the GPU box)
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
want = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__occupancy_limit_registers",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed_op_local_ld.sum", "smsp__inst_executed_op_local_st.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum"]
for r in rows[2:]:
    print("=====", r[idx["Kernel Name"]])
    for w in want:
        if w in idx:
            print(f"  {w:70s} {r[idx[w]]:>18s} {units[idx[w]]}")
    if cells:
        rd = float(r[idx["dram__bytes_read.sum"]].replace(",", "")); wr = float(r[idx["dram__bytes_write.sum"]].replace(",", ""))
        u = units[idx["dram__bytes_read.sum"]]
        mul = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}[u]
        ins = float(r[idx["smsp__inst_executed.sum"]].replace(",", ""))
        print(f"  -> DRAM bytes per cell: read {rd * mul / cells:.1f} + write {wr * mul / cells:.1f} = {(rd + wr) * mul / cells:.1f};  warp instructions per 32 cells: {ins / cells * 32:.0f}")
