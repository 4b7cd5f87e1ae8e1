"""Per-CUDA-source-line totals from `ncu --page source --csv --print-source cuda,sass`:
instructions executed, stall samples, top stall reasons.  usage: ncu_lines.py file.csv kernel_substr [top]"""
import csv, sys
path, kern, top = sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 40
rows = list(csv.reader(open(path)))
i = 0
while i < len(rows):
    r = rows[i]
    if r and r[0] == "Function Name" and kern in r[1]:
        break
    i += 1
hdr = rows[i + 1]
LN, SRC, SAMP, INST = hdr.index("Line No"), 1, hdr.index("# Samples"), hdr.index("Instructions Executed")
stall = [k for k, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
lines = []
for r in rows[i + 2:]:
    if r and r[0] == "Function Name":
        break
    if len(r) != len(hdr) or not r[LN]:
        continue          # SASS row (belongs to the line above; the line row already carries the totals)
    st = sorted(((int(r[k] or 0), hdr[k][6:]) for k in stall), reverse=True)[:3]
    lines.append((int(r[LN]), r[SRC].strip(), int(r[SAMP] or 0), int(r[INST] or 0), st))
ti, ts = sum(l[3] for l in lines), sum(l[2] for l in lines)
print(f"{kern}: instructions {ti}  samples {ts}")
print("--- by instructions")
for l in sorted(lines, key=lambda l: -l[3])[:top]:
    print(f"{l[0]:5d} inst {l[3]:9d} {100*l[3]/ti:5.1f}%  samp {100*l[2]/max(ts,1):5.1f}%  {l[1][:100]}")
print("--- by stall samples")
for l in sorted(lines, key=lambda l: -l[2])[:top]:
    print(f"{l[0]:5d} samp {l[2]:7d} {100*l[2]/max(ts,1):5.1f}%  {l[1][:80]:80s} {[(a, b) for a, b in l[4] if a]}")
