"""Summarise an .ncu-rep (raw page) into the handful of counters DESIGN.md / profiles/ quote.
usage: python tools/ncu_summary.py file.ncu-rep [more.ncu-rep ...]"""
import csv
import io
import re
import subprocess
import sys

WANT = re.compile(
    r"^(gpu__time_duration\.sum|dram__bytes_(read|write)\.sum|dram__throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"sm__warps_active\.avg\.pct_of_peak_sustained_active|smsp__inst_executed\.sum|"
    r"sm__issue_active\.avg\.pct_of_peak_sustained_elapsed|launch__registers_per_thread|launch__occupancy_limit_\w+|"
    r"launch__grid_size|launch__block_size|smsp__average_warps_issue_stalled_\w+_per_issue_active\.ratio|"
    r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum|smsp__thread_inst_executed_per_inst_executed\.ratio|"
    r"sm__inst_executed_pipe_(alu|fma|lsu)\.avg\.pct_of_peak_sustained_active|sm__pipe_fmaheavy_cycles_active\.avg\.pct_of_peak_sustained_elapsed|l1tex__t_sector_hit_rate\.pct|"
    r"lts__t_sector_hit_rate\.pct|lts__t_bytes\.sum|l1tex__t_bytes\.sum|lts__throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"l1tex__throughput\.avg\.pct_of_peak_sustained_elapsed|smsp__inst_executed_op_(shared|global)_(ld|st)\.sum|"
    r"sm__sass_inst_executed_op_(shared|global)_(ld|st)\.sum|l1tex__lsu_writeback_active\.avg\.pct_of_peak_sustained_elapsed)$")

for f in sys.argv[1:]:
    raw = subprocess.run(["ncu", "-i", f, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")]
        print("== %s :: %s" % (f, name))
        for h, u, v in zip(hdr, units, r):
            if WANT.match(h):
                if "stalled" in h and float(v or 0) < 0.2:
                    continue
                print("  %-90s %s %s" % (h, v, u))
