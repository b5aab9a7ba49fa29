#!/bin/bash
# round 2, call C: tests again, tuning variants of the snapshot and pair kernels, grid3d check
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/r2c_pytest.log
for v in "" snap_pf0 snap_pf4 snap_pf16; do
  if [ -n "$v" ]; then export BF_LIB_PATH=$PWD/baryonification_b200/variants/libbf_$v.so; else unset BF_LIB_PATH; fi
  echo "== snapshot variant [$v]"
  timeout 300 python bench.py --workload snapshot --steps 10 --warmup 3 --no-cpu --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'])"
done
for v in "" g_mb3 g_mb5 g_c96 g_c192; do
  if [ -n "$v" ]; then export BF_LIB_PATH=$PWD/baryonification_b200/variants/libbf_$v.so; else unset BF_LIB_PATH; fi
  echo "== gather variant [$v]"
  timeout 300 python bench.py --steps 5 --warmup 2 --no-cpu --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['config']['stages_ms_max_over_ranks'], d['config']['pairs'])"
done
unset BF_LIB_PATH
timeout 600 python bench.py --workload grid3d --ngrid 512 --nhalo 20000 --steps 3 --warmup 2 --no-cpu > gpurun_out/r2c_grid3d_512.json 2> gpurun_out/r2c_grid3d_512.err; echo "grid3d rc=$?"; tail -c 1500 gpurun_out/r2c_grid3d_512.json; tail -3 gpurun_out/r2c_grid3d_512.err
