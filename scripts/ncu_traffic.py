#!/usr/bin/env python
"""profiles/traffic.json from an `ncu --set full` report: per kernel, dram__bytes_read.sum + dram__bytes_write.sum of ONE
launch, stamped with the SHA-256 of the kernel's source file AS IT IS NOW (run this right after the capture, before
editing the source).  bench.py reports `roofline.traffic` from this file only while the stamp still matches.

    python scripts/ncu_traffic.py gpurun_out/r2_gather.ncu-rep bf_shell_gather.cu [--summary profiles/r02_gather_ncu.txt]
"""
import csv
import hashlib
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEEP = re.compile(r"dram__bytes_(read|write)\.sum$|dram__bytes_(read|write)\.sum\.per_second|gpu__time_duration\.sum|"
                  r"gpu__dram_throughput\.avg\.pct|launch__(registers_per_thread$|grid_size|block_size|shared_mem_per_block|"
                  r"occupancy_limit|waves)|l1tex__t_sector_hit_rate|lts__t_sector_hit_rate|l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum|"
                  r"sm__inst_executed_pipe_(alu|fma|fp64|lsu|xu|uniform)\.avg\.pct_of_peak_sustained_active|"
                  r"sm__throughput\.avg\.pct|sm__warps_active\.avg\.pct|smsp__issue_active\.avg\.pct|smsp__inst_executed\.sum$|"
                  r"smsp__thread_inst_executed_per_inst_executed\.ratio|smsp__average_warps_issue_stalled_.*_per_issue_active")


def short(name):
    m = re.match(r"(?:void )?([A-Za-z0-9_]+)(<.*?>)?\(", name)
    if not m:
        return name
    targs = (m.group(2) or "").replace("(int)", "").replace("(bool)", "").replace(" ", "")
    targs = targs.replace("1>", "1>").replace("true", "true")
    return m.group(1) + targs


def main():
    rep, source = sys.argv[1], sys.argv[2]
    summary = sys.argv[sys.argv.index("--summary") + 1] if "--summary" in sys.argv else None
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    sha = hashlib.sha256(open(os.path.join(ROOT, "baryonification_b200", "csrc", source), "rb").read()).hexdigest()[:16]
    path = os.path.join(ROOT, "profiles", "traffic.json")
    out = json.load(open(path)) if os.path.exists(path) else {}
    lines = []
    for r in rows[2:]:
        rec = dict(zip(hdr, r))
        name = short(rec.get("Kernel Name", "?"))

        def val(k):
            v, u = float(rec[k].replace(",", "")), units[hdr.index(k)].lower()
            return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12}.get(u, 1)

        nbytes = val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
        out[name] = {"bytes": nbytes, "read": val("dram__bytes_read.sum"), "write": val("dram__bytes_write.sum"),
                     "src": source, "src_sha16": sha, "report": os.path.basename(rep),
                     "duration_ms_under_ncu": float(rec["gpu__time_duration.sum"].replace(",", "")) *
                     {"ms": 1, "us": 1e-3, "ns": 1e-6, "s": 1e3}.get(units[hdr.index("gpu__time_duration.sum")].lower(), 1)}
        lines.append("Kernel: " + rec.get("Kernel Name", "?"))
        for h, u in zip(hdr, units):
            if KEEP.search(h):
                lines.append("  %-95s %-16s %s" % (h, u, rec[h]))
    json.dump(out, open(path, "w"), indent=1, sort_keys=True)
    if summary:
        open(summary, "w").write("\n".join(lines) + "\n")
    print(json.dumps({k: out[k] for k in out}, indent=1))


if __name__ == "__main__":
    main()
