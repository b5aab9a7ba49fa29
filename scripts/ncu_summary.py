"""Key metrics of every kernel in an .ncu-rep as plain text (for profiles/).  Usage: ncu_summary.py REPORT [regex]"""
import csv
import io
import re
import subprocess
import sys

KEYS = [r"^gpu__time_duration\.sum$", r"^dram__bytes_(read|write)\.sum$", r"^dram__bytes_(read|write)\.sum\.per_second$",
        r"^gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed$", r"^sm__throughput\.avg\.pct_of_peak_sustained_elapsed$",
        r"^lts__t_sector_hit_rate\.pct$", r"^l1tex__t_sector_hit_rate\.pct$", r"^lts__t_bytes\.sum$", r"^lts__t_sectors_srcunit_tex_op_read\.sum$",
        r"^sm__warps_active\.avg\.pct_of_peak_sustained_active$", r"^smsp__issue_active\.avg\.pct_of_peak_sustained_active$",
        r"^sm__inst_executed_pipe_(alu|fma|fp64|lsu|xu|fmaheavy)\.avg\.pct_of_peak_sustained_active$", r"^smsp__inst_executed\.sum$",
        r"^smsp__thread_inst_executed_per_inst_executed\.ratio$", r"^launch__(registers_per_thread|grid_size|block_size|waves_per_multiprocessor)$",
        r"^launch__occupancy_limit_(registers|shared_mem|warps|blocks)$", r"^launch__shared_mem_per_block_(static|dynamic)$",
        r"^smsp__average_warps_issue_stalled_.*_per_issue_active\.ratio$", r"^l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum$"]
rep = sys.argv[1]
pat = re.compile(sys.argv[2]) if len(sys.argv) > 2 else None
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
ki = hdr.index("Kernel Name")
for r in rows[2:]:
    if pat and not pat.search(r[ki]):
        continue
    print("Kernel:", r[ki])
    for h, u, v in zip(hdr, units, r):
        if any(re.search(k, h) for k in KEYS):
            print(f"  {h:95s} {u:18s} {v}")
