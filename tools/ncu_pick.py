"""Concise view of an ncu report's raw page: python tools/ncu_pick.py REPORT.ncu-rep [regex]"""
import csv, io, re, subprocess, sys
rep = sys.argv[1]
pat = re.compile(sys.argv[2] if len(sys.argv) > 2 else
                 r"^(gpu__time_duration.sum|dram__bytes_(read|write).sum|launch__registers_per_thread|sm__warps_active.avg.pct_of_peak_sustained_active|"
                 r"smsp__issue_active.avg.pct_of_peak_sustained_active|sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active|"
                 r"l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed|l1tex__data_pipe_lsu_wavefronts_mem_shared(_op_ld|_op_st)?.sum|"
                 r"l1tex__data_pipe_lsu_wavefronts_mem_lgds.sum|l1tex__data_bank_conflicts_pipe_lsu_mem_shared(_op_ld|_op_st)?.sum|"
                 r"smsp__average_warps_issue_stalled_.*_per_issue_active.ratio|smsp__inst_executed.sum|sm__cycles_active.avg|"
                 r"lts__t_sector_hit_rate.pct|sm__inst_executed_pipe_(lsu|alu|fma|fmaheavy|xu|uniform).avg.pct_of_peak_sustained_active|"
                 r"smsp__inst_executed_op_shared.*sum|sm__sass_inst_executed_op_shared.*sum|l1tex__throughput.avg.pct_of_peak_sustained_elapsed)$")
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[0]
ki = hdr.index("Kernel Name")
for r in rows[2:]:
    print("==", r[ki][:80])
    for h, v in zip(hdr, r):
        if pat.search(h):
            try:
                fv = float(v)
                if "stalled" in h and fv < 0.3:
                    continue
            except ValueError:
                pass
            print(f"   {h:100s} {v}")
