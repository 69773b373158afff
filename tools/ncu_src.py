"""Top source lines by warp-stall samples from `ncu --page source --csv --print-source sass,cuda`:
python tools/ncu_src.py REPORT.ncu-rep KERNEL_REGEX [N]"""
import csv, io, subprocess, sys, collections
rep, kern = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda", "-k", "regex:" + kern],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
hdr = rows[hi]
ci = {h: i for i, h in enumerate(hdr)}
line_i, samp_i, inst_i = 0, hdr.index("# Samples"), hdr.index("Instructions Executed")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
agg = collections.defaultdict(lambda: [0, 0, collections.Counter(), ""])
tot = 0
for r in rows[hi + 1:]:
    if len(r) < len(hdr):
        continue
    try:
        s = int(r[samp_i] or 0)
    except ValueError:
        continue
    key = r[line_i]
    a = agg[key]
    a[0] += s; tot += s
    a[1] += int(r[inst_i] or 0)
    a[3] = r[1][:110]
    for i in stall_cols:
        try:
            v = int(r[i] or 0)
        except ValueError:
            v = 0
        if v:
            a[2][hdr[i]] += v
print("total samples", tot)
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:n]:
    top = ", ".join(f"{k[6:]}={v}" for k, v in a[2].most_common(3))
    print(f"{a[0]:7d} {100.0 * a[0] / max(tot, 1):5.1f}%  inst {a[1]:9d}  L{key:>5s}  {top:45s} | {a[3].strip()}")
