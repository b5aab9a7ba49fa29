#!/bin/bash
# quick 2-GPU check of the paint / shell end-to-end paths
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
for w in paint shell; do
timeout 600 $TR bench.py --gpus 2 --workload $w --steps 5 --warmup 3 --no-cpu > gpurun_out/r2e_${w}_n2.json 2> gpurun_out/r2e_${w}_n2.err; echo "bench $w rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2e_${w}_n2.json')); print('$w n2', d['ms_per_step'], d['e2e']['ms'], d['e2e'].get('stages_ms'), d['config'].get('n1_equal'))"
done
