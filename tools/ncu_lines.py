"""Aggregate an `ncu --page source --print-source cuda,sass --csv` dump per CUDA source line."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 50
fname = None; hdr = None; data = []
for r in rows:
    if len(r) == 2 and r[0] == "File Path": fname = r[1].split('/')[-1]; continue
    if len(r) == 2 and r[0] == "Function Name": continue
    if len(r) > 3 and r[0] == "Line No": hdr = r; iI = hdr.index("Instructions Executed"); iS = hdr.index("# Samples"); continue
    if hdr and len(r) > iI and r[2] == "-":   # a CUDA source line row (address '-')
        try: data.append((int(r[iI]), int(r[iS]), fname, r[0], r[1]))
        except ValueError: pass
tot = sum(d[0] for d in data); ts = sum(d[1] for d in data)
print("total inst", tot, "samples", ts)
for d in sorted(data, key=lambda d: -d[0])[:top]:
    print(f"{100*d[0]/tot:5.2f}% inst {100*d[1]/max(ts,1):5.2f}% smp {d[2]}:{d[3]:>4}  {d[4][:100]}")
