"""Summarise `ncu --page source --csv` output: top SASS lines by stall samples."""
import csv
import sys

path, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25
rows = list(csv.reader(open(path)))
hdr = rows[1]
si, ai, ni = hdr.index("Source"), hdr.index("Address"), hdr.index("# Samples")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
body = []
for r in rows[2:]:
    if r and r[0] == "Kernel Name":
        break                      # first kernel instance only
    if len(r) == len(hdr) and r[ni].isdigit():
        body.append(r)
tot = sum(int(r[ni] or 0) for r in body)
print("total samples", tot, "instructions", len(body))
agg = {}
for r in body:
    for i in stall_cols:
        agg[hdr[i]] = agg.get(hdr[i], 0) + int(r[i] or 0)
print("stall mix:", {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]})
for r in sorted(body, key=lambda r: -int(r[ni] or 0))[:top]:
    st = sorted(((int(r[i] or 0), hdr[i]) for i in stall_cols), reverse=True)[:2]
    print(f"{int(r[ni]):6d} {100*int(r[ni])/max(tot,1):5.1f}%  {r[si][:90]:90s} {st}")
