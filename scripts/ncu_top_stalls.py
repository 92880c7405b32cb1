"""Top warp-stall sampling lines of an ncu report's source page (SASS view), with the dominant stall reason of each.
usage: python scripts/ncu_top_stalls.py report.ncu-rep [N]"""
import csv
import subprocess
import sys

rep, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 20
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
print(rows[0][1][:120])
hdr = rows[1]
ci = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") or h.startswith("Stall")]
data = []
for r in rows[2:]:
    try:
        n = int(r[ci["# Samples"]])
    except (ValueError, IndexError, KeyError):
        continue
    reasons = []
    for h in stall_cols:
        try:
            v = int(r[ci[h]])
        except ValueError:
            continue
        if v:
            reasons.append((v, h))
    reasons.sort(reverse=True)
    data.append((n, r[ci["Source"]].strip()[:90], reasons[:2]))
tot = sum(d[0] for d in data) or 1
print("total samples", tot, "| stall columns:", len(stall_cols))
for n, src, reasons in sorted(data, key=lambda d: -d[0])[:top]:
    print(f"{n:7d} {100 * n / tot:5.1f}%  {src:90s} {reasons}")
