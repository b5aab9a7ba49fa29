#!/bin/bash
# round 2, call A: all GPU tests, default bench, ncu capture of the pair kernel
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/r2a_smi.txt; nproc >> gpurun_out/r2a_smi.txt; free -g >> gpurun_out/r2a_smi.txt
timeout 1200 python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?"
tail -30 gpurun_out/r2a_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"
cat gpurun_out/r2a_bench.json
timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e > gpurun_out/r2a_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gather_kernel -c 1 -o gpurun_out/r2a_gather \
   python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e > gpurun_out/r2a_ncu.log 2>&1; echo "ncu rc=$?"
