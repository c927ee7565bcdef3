"""Per-source-line share of executed warp instructions / stall samples from an .ncu-rep (needs -lineinfo + --import-source)."""
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur = None
agg = {}
for r in csv.reader(out.splitlines()):
    if not r:
        continue
    if r[0] == 'File Path':
        cur = r[1].split('/')[-1]
        continue
    if r[0].isdigit() and len(r) > 7:
        inst = int(r[7]) if r[7].isdigit() else 0
        smp = int(r[6]) if r[6].isdigit() else 0
        a = agg.setdefault((cur, int(r[0])), [0, 0, r[1].strip()[:100]])
        a[0] += inst
        a[1] += smp
tot = sum(v[0] for v in agg.values()) or 1
ts = sum(v[1] for v in agg.values()) or 1
print(f"total warp instructions {tot}, stall samples {ts}")
for (f, l), (i, s, src) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{100 * s / ts:5.1f}% smp {100 * i / tot:5.1f}% inst  {f}:{l}  {src}")
