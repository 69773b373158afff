"""Summarise ncu output for profiles/.
  python tools/ncu_summary.py launches <launches.csv>          -> per-kernel totals / shares (gpu__time_duration)
  python tools/ncu_summary.py full <report.ncu-rep>            -> key counters per captured kernel (needs ncu)"""
import csv, subprocess, sys, io, collections

def launches(path):
    rows = [r for r in csv.reader(open(path, errors="ignore")) if r and not r[0].startswith("==")]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        if len(r) <= vi: continue
        v = float(r[vi].replace(",", ""))
        u = r[ui]
        us = v / 1e3 if u in ("ns", "nsecond") else v * 1e3 if u in ("ms", "msecond") else v
        n = r[ki].split("(")[0][:90].replace(",", ";")
        a = agg.setdefault(n, [0, 0.0]); a[0] += 1; a[1] += us
    tot = sum(a[1] for a in agg.values())
    print(f"# total device time {tot/1e3:.3f} ms over {sum(a[0] for a in agg.values())} launches")
    print("kernel,launches,total_us,avg_us,share")
    for n, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{n},{a[0]},{a[1]:.1f},{a[1]/a[0]:.2f},{a[1]/tot:.4f}")

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "lts__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum",
        "sm__cycles_active.avg", "sm__cycles_active.max"]

def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    print("kernel," + ",".join(f"{w} [{units[hdr.index(w)]}]" for w in WANT if w in hdr))
    for r in rows[2:]:
        print(r[ki].split("(")[0][-60:].replace(",", ";") + "," + ",".join(r[hdr.index(w)] for w in WANT if w in hdr))

if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
