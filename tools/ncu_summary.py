"""Per-launch summary of an .ncu-rep: python tools/ncu_summary.py rep.ncu-rep"""
import csv, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines())); hdr = rows[0]
want = {"gpu__time_duration.sum": "us", "dram__bytes_read.sum": "rd", "dram__bytes_write.sum": "wr",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed": "dram%", "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue%",
        # tensor pipe: the collected counter over the SM's ACTIVE cycles (at whatever clock the kernel ran) -- the TriageCompute
        # "realtime" figure of the same pipe is sampled and varies 2x between replays of one kernel, it is printed only for reference
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_active%",
        "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed": "tensor_rt%",
        "sm__cycles_active.avg": "sm_cycles",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed": "lts%", "launch__registers_per_thread": "regs"}
src = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout.splitlines()
blocks, cur = [], None
for line in src:
    if line.startswith('"Kernel Name"'):
        cur = []; blocks.append(cur)
    elif cur is not None:
        cur.append(line)
blocks = blocks[::2] if len(blocks) == 2 * (len(rows) - 2) else blocks
for n, r in enumerate(rows[2:]):
    name = r[hdr.index("Kernel Name")][:48]
    vals = " ".join(f"{lab}={float(r[hdr.index(k)]):.4g}{rows[1][hdr.index(k)] if lab in ('us','rd','wr') else ''}" for k, lab in want.items() if k in hdr and r[hdr.index(k)] not in ("", "no data", "n/a"))
    st = ""
    if n < len(blocks):
        b = list(csv.reader(blocks[n])); h = b[0]; b = b[1:]
        cols = [i for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
        agg = {h[i][6:]: sum(int(x[i] or 0) for x in b) for i in cols}
        tot = sum(agg.values()) or 1
        st = f" sass={len(b)} " + " ".join(f"{k}={v / tot:.2f}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:7])
    print(f"{n:2d} {name:48s} {vals}{st}")
