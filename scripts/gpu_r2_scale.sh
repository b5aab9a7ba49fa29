#!/bin/bash
# round 2: the four bench workloads on N GPUs of one box (N = $1), optionally the N-GPU == 1-GPU functional check first
N=${1:-2}
CHECK=${2:-1}
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8 > $O/r2s_n${N}_smi.txt
if [ "$CHECK" = "1" ]; then
  timeout 900 $TR scripts/check_sharded.py 1024 50000 > $O/r2s_check_n${N}.log 2>&1; echo "check rc=$?"; grep -E "world|SHARDED|rror" $O/r2s_check_n${N}.log | tail -20
fi
for w in shell paint snapshot; do
  timeout 900 $TR bench.py --gpus $N --workload $w --steps 10 --warmup 3 --no-cpu > $O/r2s_bench_${w}_n${N}.json 2> $O/r2s_bench_${w}_n${N}.err; echo "bench $w rc=$?"
  grep -v "^W\|^\[W\|NCCL version\|^$" $O/r2s_bench_${w}_n${N}.err | tail -4
done
timeout 900 $TR bench.py --gpus $N --workload grid3d --steps 3 --warmup 2 --no-cpu > $O/r2s_bench_grid3d_n${N}.json 2> $O/r2s_bench_grid3d_n${N}.err; echo "grid3d rc=$?"
grep -v "^W\|^\[W\|NCCL version\|^$" $O/r2s_bench_grid3d_n${N}.err | tail -4
python - $N <<'PY'
import json, sys
N = sys.argv[1]
for w in ("shell", "paint", "snapshot", "grid3d"):
    try:
        d = json.load(open(f"gpurun_out/r2s_bench_{w}_n{N}.json"))
        c = d["config"]
        print(w, "N", d["n_gpus"], round(d["ms_per_step"], 2), "ms/step", "%.4g" % d["value"], d["unit"], "roofline", round(d["roofline"]["frac"], 3),
              "e2e ms", d["e2e"] and round(d["e2e"]["ms"], 1), c.get("stages_ms_max_over_ranks"), "n1_equal", c.get("n1_equal"), c.get("n1_max_abs_diff_over_scale"), "clk", d["clocks"])
    except Exception as e:
        print(w, "failed", e)
PY
