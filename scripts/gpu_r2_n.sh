#!/bin/bash
# refresh of the N = 1 lines whose end-to-end path changed (uploads of page-locked inputs start after the set-up)
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python bench.py --steps 20 --warmup 5 > $O/r2n_bench_shell_n1.json 2> $O/r2n_bench_shell_n1.err; echo "bench shell rc=$?"
timeout 600 python bench.py --workload paint --steps 10 --warmup 3 > $O/r2n_bench_paint_n1.json 2> $O/r2n_bench_paint_n1.err; echo "bench paint rc=$?"
timeout 600 python bench.py --workload snapshot --steps 10 --warmup 3 > $O/r2n_bench_snapshot_n1.json 2> $O/r2n_bench_snapshot_n1.err; echo "bench snapshot rc=$?"
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "snapshot or page_locked or shell_parity or tiny" 2>&1 | tail -2
python - <<'PY'
import json
for w in ("shell", "paint", "snapshot"):
    d = json.load(open(f"gpurun_out/r2n_bench_{w}_n1.json"))
    print(w, round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["ms"], 1), d["e2e"].get("stages_ms"), "roofline", round(d["roofline"]["frac"], 3), d["roofline"].get("dram"))
PY
