#!/bin/bash
# pair kernel with the one-word cell check: parity (shell tests, NSIDE 4096 sub-catalogue, C2) and timing
mkdir -p gpurun_out
timeout 45 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e 2>/dev/null > gpurun_out/r2t_bench_shell.json; python -c "
import json; d=json.load(open('gpurun_out/r2t_bench_shell.json')); print('shell', d['ms_per_step'], d['config']['stages_ms_max_over_ranks'], d['config']['pairs'], d['config']['mass_conservation_relerr'])"
timeout 45 python bench.py --workload paint --steps 5 --warmup 3 --no-cpu --no-e2e 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('paint', d['ms_per_step'], d['config']['pairs'])"
timeout 100 python -m pytest tests/test_gpu_parity.py tests/test_gpu_baseline_sizes.py -m gpu -x -q -k "shell or headline or c2 or paint" 2>&1 | tail -2
