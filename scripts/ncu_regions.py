"""Warp-stall samples of one kernel in an .ncu-rep, aggregated over barrier-delimited SASS regions (development aid).
usage: python scripts/ncu_regions.py report.ncu-rep [min_share]"""
import csv, io, subprocess, sys

rep = sys.argv[1]
min_share = float(sys.argv[2]) if len(sys.argv) > 2 else 0.01
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
h = rows[1]
ia, isrc, isamp, iex = h.index("Address"), h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
regs, cur, tot = [], dict(start=None, samp=0, n=0, ops={}, ex=0), 0
for r in rows[2:]:
    if len(r) <= isamp:
        continue
    try:
        s = int(r[isamp])
    except ValueError:
        s = 0
    try:
        ex = int(r[iex])
    except ValueError:
        ex = 0
    toks = r[isrc].split()
    op = toks[1] if toks and toks[0].startswith("@") and len(toks) > 1 else (toks[0] if toks else "")
    cur["samp"] += s; cur["n"] += 1; cur["ex"] += ex; tot += s
    k = op.split(".")[0]
    cur["ops"][k] = cur["ops"].get(k, 0) + 1
    if cur["start"] is None:
        cur["start"] = r[ia]
    if op.startswith("BAR"):
        regs.append(cur); cur = dict(start=None, samp=0, n=0, ops={}, ex=0)
regs.append(cur)
print("total samples", tot, "regions", len(regs))
for i, g in enumerate(regs):
    if g["samp"] > tot * min_share:
        ops = sorted(g["ops"].items(), key=lambda x: -x[1])[:8]
        print(i, g["start"][-6:], "share %.3f" % (g["samp"] / tot), "ninst", g["n"], "exec", g["ex"], ops)
