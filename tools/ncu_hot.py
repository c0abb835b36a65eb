"""Hot SASS lines of one kernel from an .ncu-rep (source page): python tools/ncu_hot.py rep.ncu-rep 'k_sim_tc<(int)2' [top]"""
import csv, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout.splitlines()
blocks, cur = [], None
for line in out:
    if line.startswith('"Kernel Name"'):
        cur = {"name": line, "lines": []}; blocks.append(cur)
    elif cur is not None:
        cur["lines"].append(line)
for blk in blocks:
    if pat not in blk["name"]:
        continue
    rows = list(csv.reader(blk["lines"]))
    hdr = rows[0]; rows = rows[1:]
    ia, isrc, isamp, iex = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(int(r[isamp] or 0) for r in rows)
    print(blk["name"][:120], "total samples", tot, "instructions", len(rows))
    agg = {}
    for r in rows:
        for i in stall_cols:
            agg[hdr[i]] = agg.get(hdr[i], 0) + int(r[i] or 0)
    print({k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
    order = sorted(range(len(rows)), key=lambda i: -int(rows[i][isamp] or 0))[:top]
    for i in sorted(order):
        r = rows[i]
        st = {hdr[c][6:]: int(r[c]) for c in stall_cols if r[c] and int(r[c]) > 0}
        st = dict(sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print(f"{i:5d} {r[isamp]:>6} ex={r[iex]:>9} {r[isrc][:70]:70s} {st}")
