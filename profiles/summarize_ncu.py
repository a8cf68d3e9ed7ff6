"""Prints the per-kernel summary of an `ncu --page raw --csv` export (the numbers quoted in profiles/*.md)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]
idx = {h: i for i, h in enumerate(hdr)}
want = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum"]
stall = [h for h in hdr if "pcsamp_warps_issue_stalled" in h and not h.endswith("_not_issued")]
for r in rows[2:]:
    print("=====", r[idx["Kernel Name"]][:70])
    for k in want:
        if k in idx:
            print("   ", k, r[idx[k]], rows[1][idx[k]])
    st = []
    for k in stall:
        try:
            st.append((float(r[idx[k]].replace(",", "")), k))
        except ValueError:
            pass
    tot = sum(v for v, _ in st) or 1
    print("    stalls:", ", ".join(f"{k.split('stalled_')[1]} {100*v/tot:.0f}%" for v, k in sorted(st, reverse=True)[:7]))
