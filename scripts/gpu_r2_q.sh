#!/bin/bash
# last call of the round: smoke(), the tests touched by the last host-side changes, the no-alloc test alone
mkdir -p gpurun_out
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 100 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -1
timeout 150 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "paint or band_shard or page_locked or make_map or snapshot_parity" 2>&1 | tail -2
timeout 90 python bench.py --workload paint --steps 3 --warmup 3 --no-cpu 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('paint n1', d['ms_per_step'], d['e2e']['ms'], d['e2e']['stages_ms'])"
