#!/bin/bash
# round 2, call I: TMA-fed streaming snapshot kernel -- parity tests, memcheck of a small case, timing against the chunked
# kernel and ring-depth / occupancy variants
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "snapshot or delta_pk" > $O/r2i_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/r2i_pytest.log
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "snapshot_parity and reference" > $O/r2i_memcheck.log 2>&1; echo "memcheck rc=$?"; tail -4 $O/r2i_memcheck.log
timeout 600 compute-sanitizer --tool racecheck --error-exitcode 9 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "snapshot_parity and reference and 3-64" > $O/r2i_racecheck.log 2>&1; echo "racecheck rc=$?"; tail -4 $O/r2i_racecheck.log
run() { timeout 300 python bench.py --workload snapshot --steps 10 --warmup 3 --no-cpu --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['config']['pairs'])"; }
echo "== stream (default)"; run
echo "== chunked"; BF_SNAP_CHUNKED=1 run
for v in ss_d4 ss_d4_mb6 ss_d6_mb4; do
  echo "== variant $v"; BF_LIB_PATH=$PWD/baryonification_b200/variants/libbf_$v.so run
done
