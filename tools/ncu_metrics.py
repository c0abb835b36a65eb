"""Selected raw metrics of every launch in an .ncu-rep: python tools/ncu_metrics.py rep.ncu-rep [substring ...]
Default selection: pipe utilisation, shared-memory wavefronts, issue, stalls-independent counters."""
import csv, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines())); hdr = rows[0]; units = rows[1]
pats = sys.argv[2:] or ["gpu__time_duration.sum", "sm__inst_executed_pipe_", "sm__pipe_", "smsp__issue_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared",
                        "l1tex__data_bank_", "smsp__inst_executed.sum", "sm__throughput", "l1tex__throughput", "lts__throughput", "dram__throughput",
                        "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__warps_active", "smsp__cycles_active.avg", "l1tex__lsu_writeback", "launch__registers"]
for r in rows[2:]:
    print("==", r[hdr.index("Kernel Name")][:80])
    for i, h in enumerate(hdr):
        if any(p in h for p in pats) and r[i] not in ("", "0"):
            print(f"   {h} = {r[i]} {units[i]}")
