#!/bin/bash
# round 2, call L (2 GPUs): make_map tests, sharded paint with pipelined read-back
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "make_map or paint or page_locked" 2>&1 | tail -4
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
timeout 900 $TR scripts/check_sharded.py 1024 50000 > gpurun_out/r2l_check_n2.log 2>&1; echo "check rc=$?"; grep -E "Shell|SHARDED|rror" gpurun_out/r2l_check_n2.log | tail -6
for w in paint; do
timeout 600 $TR bench.py --gpus 2 --workload $w --steps 5 --warmup 3 --no-cpu > gpurun_out/r2e_${w}_n2.json 2> gpurun_out/r2e_${w}_n2.err; echo "bench $w rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2e_${w}_n2.json')); print('$w n2', d['ms_per_step'], d['e2e']['ms'], d['e2e'].get('stages_ms'), d['config'].get('n1_equal'))"
done
timeout 600 python bench.py --workload paint --steps 5 --warmup 3 --no-cpu > gpurun_out/r2e_paint_n1.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/r2e_paint_n1.json')); print('paint n1', d['ms_per_step'], d['e2e']['ms'], d['e2e'].get('stages_ms'))"
