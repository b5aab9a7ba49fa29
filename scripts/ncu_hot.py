"""Hottest CUDA source lines of the profiled kernel(s) by warp-stall samples.  Usage: ncu_hot.py REPORT [N]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
cur, hdr, data = None, None, []
def f(x):
    try: return float(x)
    except Exception: return 0.0
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1]; continue
    if len(r) >= 2 and r[0] == "Line No":
        hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0] != "":
        data.append((cur, r))
iS = hdr.index("# Samples"); iI = hdr.index("Instructions Executed")
st = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(f(r[iS]) for _, r in data) or 1
data.sort(key=lambda d: -f(d[1][iS]))
print(f"total samples {tot:.0f}")
for cur, r in data[:n]:
    top = sorted(((f(r[i]), h) for i, h in st), reverse=True)[:2]
    print(f"{100*f(r[iS])/tot:5.1f}%  {cur.split('/')[-1]}:{r[0]:>4s} inst={r[iI]:>9s} "
          f"{' '.join(f'{h[6:]}={int(v)}' for v, h in top):32s}| {r[1].strip()[:100]}")
