"""Print selected metrics of every launch in an .ncu-rep: python tools/ncu_summary.py file.ncu-rep"""
import csv, subprocess, sys, io
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
want = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "us"), ("launch__registers_per_thread", "regs"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("smsp__thread_inst_executed_per_inst_executed.ratio", "thr/inst"),
        ("smsp__inst_executed.sum", "winst"), ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("l1tex__t_sector_hit_rate.pct", "L1hit%"), ("lts__t_sector_hit_rate.pct", "L2hit%"),
        ("dram__bytes_read.sum", "dramR"), ("dram__bytes_write.sum", "dramW"), ("lts__t_bytes.sum", "L2bytes"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "L1tp%"), ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2tp%"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "st_long"),
        ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "st_wait"),
        ("smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "st_br"),
        ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "st_lg"),
        ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "st_short"),
        ("smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "st_notsel"),
        ("smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "st_noinst"),
        ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "st_math"),
        ("smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "st_mio")]
idx = [(hdr.index(k), n) for k, n in want if k in hdr]
for r in rows[2:]:
    parts = []
    for i, n in idx:
        v = r[i]
        if n == "kernel":
            v = v.split("(")[0].replace("ptb::", "").replace("void ", "")
        else:
            try:
                v = "%.3g" % float(v.replace(",", ""))
            except ValueError:
                pass
            v = v + (units[i] if n in ("dramR", "dramW", "L2bytes") else "")
        parts.append("%s=%s" % (n, v))
    print(" ".join(parts))
