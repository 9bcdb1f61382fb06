"""Top stall locations of an ncu report's SASS source page.
    python scripts/ncu_hot.py report.ncu-rep [kernel index] [top N]
"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks = out.split('"Kernel Name",')
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
blk = blocks[1 + which]
lines = blk.split("\n")
print("kernel:", lines[0][:100])
rows = list(csv.reader(io.StringIO("\n".join(lines[1:]))))
hdr = rows[0]
idx = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = []
total = 0
for i, r in enumerate(rows[1:]):
    if len(r) < len(hdr):
        continue
    n = int(r[idx["# Samples"]] or 0)
    total += n
    data.append((n, i, r))
print("total samples", total)
agg = {s: sum(int(r[idx[s]] or 0) for _, _, r in data) for s in stalls}
print("by reason:", ", ".join(f"{k[6:]} {v * 100 // max(total, 1)}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v * 100 // max(total, 1) > 0))
for n, i, r in sorted(data, key=lambda t: -t[0])[:top]:
    why = sorted(((int(r[idx[s]] or 0), s[6:]) for s in stalls), reverse=True)[:2]
    print(f"{n * 100 / max(total, 1):5.1f}%  #{i:4d} exec {r[idx['Instructions Executed']]:>9}  {r[idx['Source']].strip()[:70]:70s} {why[0][1]}:{why[0][0]} {why[1][1]}:{why[1][0]}")
