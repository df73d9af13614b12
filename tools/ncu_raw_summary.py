"""Key metrics per kernel instance from an `ncu --set full` report.  Usage: python tools/ncu_raw_summary.py X.ncu-rep [max_per_kernel]"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]
cap = int(sys.argv[2]) if len(sys.argv) > 2 else 2
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
WANT = ["launch__grid_size", "launch__block_size", "gpu__time_duration.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "smsp__inst_executed.sum", "sm__cycles_active.avg", "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
seen = collections.Counter()
for r in rows[2:]:
    name = r[idx["Kernel Name"]]
    key = (name, r[idx["launch__grid_size"]])
    seen[key] += 1
    if seen[key] > cap:
        continue
    print("==", name)
    for w in WANT:
        if w in idx:
            print(f"   {w} = {r[idx[w]]} {units[idx[w]]}")
