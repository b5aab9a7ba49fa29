#!/bin/bash
# round 2, call K: full GPU suite after the upload refactor; two more occupancy points of the streaming snapshot kernel
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > $O/r2k_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/r2k_pytest.log
run() { timeout 300 python bench.py --workload snapshot --steps 10 --warmup 3 --no-cpu --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['config']['pairs'])"; }
for v in ss_mb5_d3 ss_mb4_d4_wq32; do
  echo "== variant $v"; BF_LIB_PATH=$PWD/baryonification_b200/variants/libbf_$v.so run
done
