#!/bin/bash
# round 2, call J: streaming snapshot kernel -- parity, variants, halo-grid resolution
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "snapshot" > $O/r2j_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r2j_pytest.log
run() { timeout 300 python bench.py --workload snapshot --steps 10 --warmup 3 --no-cpu --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['config']['pairs'])"; }
echo "== stream (default)"; run
for v in $VARIANTS; do
  echo "== variant $v"; BF_LIB_PATH=$PWD/baryonification_b200/variants/libbf_$v.so run
done
for g in 224 192 160 128 96 64; do echo "== ng $g"; BF_SNAP_NG=$g run; done
echo "== chunked ng 128"; BF_SNAP_CHUNKED=1 BF_SNAP_NG=128 run
