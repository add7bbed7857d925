"""Aggregate an `ncu --page source --print-source cuda,sass --csv` export per CUDA source line:
share of executed warp instructions and of stall samples.  usage: ncu_lines.py file.csv [min_pct]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
hdr = next(r for r in rows if r and r[0] == "Line No")
ix = {}
for i, h in enumerate(hdr):
    ix.setdefault(h, i)
agg = {}
seen = set()
cur = None


def num(x):
    try:
        return float(x)
    except ValueError:
        return 0.0


# SASS rows follow the CUDA line they belong to; inlined code is listed under several lines, so
# every SASS address is counted once (first owner wins)
for r in rows:
    if len(r) < len(hdr):
        continue
    if r[0].isdigit() and r[2] == "-":
        cur = (int(r[0]), r[1]); continue
    if not r[2].startswith("0x") or r[2] in seen or cur is None:
        continue
    seen.add(r[2])
    a = agg.setdefault(cur[0], [cur[1], 0.0, 0.0, 0.0])
    a[1] += num(r[ix["Instructions Executed"]]); a[2] += num(r[ix["# Samples"]])
    a[3] += num(r[ix["Thread Instructions Executed"]])
ti = sum(a[1] for a in agg.values()); ts = sum(a[2] for a in agg.values())
print(f"total warp instr {ti:.0f}  samples {ts:.0f}")
for ln in sorted(agg):
    s, i, m, t = agg[ln]
    if 100 * i / ti >= thr or 100 * m / ts >= thr:
        print(f"{ln:4d} {100*i/ti:5.1f}% inst {100*m/ts:5.1f}% samp  thr/inst {t/max(i,1):4.1f} | {s.strip()[:100]}")
