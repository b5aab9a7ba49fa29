#!/bin/bash
# quick 2-GPU check of the shell end-to-end path (N = 2 and N = 1)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
for w in shell paint; do
timeout 600 $TR bench.py --gpus 2 --workload $w --steps 5 --warmup 3 --no-cpu > gpurun_out/r2e_${w}_n2.json 2> gpurun_out/r2e_${w}_n2.err; echo "bench $w rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2e_${w}_n2.json')); print('$w n2', d['ms_per_step'], d['e2e']['ms'], d['e2e'].get('stages_ms'), d['config'].get('n1_equal'))"
timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu > gpurun_out/r2e_${w}_n1.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/r2e_${w}_n1.json')); print('$w n1', d['ms_per_step'], d['e2e']['ms'], d['e2e'].get('stages_ms'))"
done
timeout 600 python bench.py --workload snapshot --steps 5 --warmup 3 --no-cpu > gpurun_out/r2e_snapshot_n1.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/r2e_snapshot_n1.json')); print('snapshot n1', d['ms_per_step'], d['e2e']['ms'])"
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "page_locked or snapshot or shell or grid" 2>&1 | tail -3
