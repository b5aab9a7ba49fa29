#!/bin/bash
# final code on 8 GPUs: shell and snapshot (device step + end to end)
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517"
for w in shell snapshot; do
  timeout 400 $TR bench.py --gpus 8 --workload $w --steps 10 --warmup 3 --no-cpu > $O/r2p_bench_${w}_n8.json 2> $O/r2p_bench_${w}_n8.err; echo "bench $w rc=$?"
done
python - <<'PY'
import json
for w in ("shell", "snapshot"):
    try:
        d = json.load(open(f"gpurun_out/r2p_bench_{w}_n8.json")); c = d["config"]
        print(w, round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["ms"], 1), d["e2e"].get("stages_ms"), "n1", c.get("n1_equal"))
    except Exception as e:
        print(w, "failed", e)
PY
