"""Aggregates `ncu -i X.ncu-rep --page source --csv --print-source sass,cuda` by file and source line."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur = None
tot = {}
lines = []
for r in rows:
    if r and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if not r or r[0] in ("Function Name", "Line No", ""):
        continue
    try:
        ie, smp = int(r[7]), int(r[6])
    except (ValueError, IndexError):
        continue
    tot.setdefault(cur, [0, 0])
    tot[cur][0] += ie
    tot[cur][1] += smp
    lines.append((ie, smp, cur, r[0], r[1][:100]))
T = sum(v[0] for v in tot.values())
S = sum(v[1] for v in tot.values())
print(f"total warp instructions {T}, stall samples {S}")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1][0]):
    print(f"{k:28s} inst {v[0] / T * 100:5.1f}%  samples {v[1] / S * 100:5.1f}%")
lines.sort(reverse=True)
for ie, smp, f, ln, src in lines[:top]:
    print(f"{ie / T * 100:5.2f}% {smp / S * 100:5.2f}% {f}:{ln} {src}")
