#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r2g_noalloc_alone.log 2>&1; echo "no-alloc alone rc=$?"; tail -3 gpurun_out/r2g_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2g_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2g_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2g_bench_shell.json 2> gpurun_out/r2g_bench_shell.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2g_bench_shell.json')); print(d['ms_per_step'], d['config']['stages_ms_max_over_ranks'], d['config']['regrid_ms'], d['roofline']['fp64']); print(d['e2e']); print(d['cpu_baseline'])"
