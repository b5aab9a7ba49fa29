#!/bin/bash
# 2-GPU sanity run of the final code: shell, snapshot, grid3d (with its new N-GPU == 1-GPU check)
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
for w in shell snapshot; do
  timeout 900 $TR bench.py --gpus 2 --workload $w --steps 10 --warmup 3 --no-cpu > $O/r2o_bench_${w}_n2.json 2> $O/r2o_bench_${w}_n2.err; echo "bench $w rc=$?"
done
timeout 900 $TR bench.py --gpus 2 --workload grid3d --steps 3 --warmup 2 --no-cpu > $O/r2o_bench_grid3d_n2.json 2> $O/r2o_bench_grid3d_n2.err; echo "grid3d rc=$?"
python - <<'PY'
import json
for w in ("shell", "snapshot", "grid3d"):
    try:
        d = json.load(open(f"gpurun_out/r2o_bench_{w}_n2.json")); c = d["config"]
        print(w, round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["ms"], 1), d["e2e"].get("stages_ms"), "n1", c.get("n1_equal"), c.get("n1_error"))
    except Exception as e:
        print(w, "failed", e)
PY
