#!/bin/bash
# round 2, call B: GPU tests, bench of the four workloads at N=1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --durations=12 > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?"
tail -40 gpurun_out/r2b_pytest.log
for w in shell paint snapshot; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 > gpurun_out/r2b_bench_$w.json 2> gpurun_out/r2b_bench_$w.err; echo "bench $w rc=$?"
  tail -c 3000 gpurun_out/r2b_bench_$w.json; tail -5 gpurun_out/r2b_bench_$w.err
done
timeout 900 python bench.py --workload grid3d --steps 3 --warmup 2 > gpurun_out/r2b_bench_grid3d.json 2> gpurun_out/r2b_bench_grid3d.err; echo "bench grid3d rc=$?"
tail -c 3000 gpurun_out/r2b_bench_grid3d.json; tail -5 gpurun_out/r2b_bench_grid3d.err
