"""Aggregate an `ncu --page source --print-source cuda,sass --csv` dump by CUDA source line.
usage: python tools/ncu_lines.py dump.csv [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur = None; hdr = None; out = []
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; continue
    if r[0] == "Function Name" or hdr is None: continue
    if len(r) > 7 and r[2] == "-":       # a CUDA source line row (aggregated)
        try: out.append((int(r[4] or 0), int(r[7] or 0), cur, r[0], r[1].strip()[:120]))
        except ValueError: pass
tot = sum(o[0] for o in out) or 1
ti = sum(o[1] for o in out) or 1
print("total samples", tot, "total warp instr", ti)
for o in sorted(out, reverse=True)[:top]:
    print(f"{100*o[0]/tot:5.1f}% stall  {100*o[1]/ti:5.1f}% inst  {o[2]}:{o[3]}: {o[4]}")
