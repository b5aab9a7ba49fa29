#!/bin/bash
# round 2, call E: new tests, e2e after the set-up-in-runs change, NP=2 variants of the pair kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/r2e_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r2e_bench_shell.json 2> gpurun_out/r2e_bench_shell.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2e_bench_shell.json')); print(d['ms_per_step'], d['config']['stages_ms_max_over_ranks'], d['config']['regrid_ms']); print(d['e2e'])"
timeout 600 python bench.py --workload paint --steps 10 --warmup 3 --no-cpu 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('paint', d['ms_per_step'], d['e2e']['ms'], d['e2e']['stages_ms'])"
for v in np2_mb3 np2_mb4; do
  export BF_LIB_PATH=$PWD/baryonification_b200/variants/libbf_$v.so
  echo "== gather variant [$v]"
  timeout 300 python bench.py --steps 5 --warmup 2 --no-cpu --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['config']['stages_ms_max_over_ranks'], d['config']['pairs'])"
  timeout 300 python bench.py --workload paint --steps 5 --warmup 2 --no-cpu --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('paint', d['ms_per_step'], d['config']['stages_ms_max_over_ranks'], d['config']['pairs'])"
done
