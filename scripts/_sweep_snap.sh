for v in "-DSN_MINBLOCKS=3 -DSN_TILES=16" "-DSN_MINBLOCKS=4 -DSN_TILES=16" "-DSN_MINBLOCKS=3 -DSN_TILES=32" "-DSN_MINBLOCKS=3 -DSN_TILES=8" "-DSN_MINBLOCKS=2 -DSN_TILES=16"; do
  echo "== $v"; BF_NVCC_EXTRA="$v" python -m baryonification_b200.build > /tmp/b.log 2>&1; grep -A3 "snap_gather_kernelILi3ELb1" /tmp/b.log | grep -E "Used|spill" | tr '\n' ' '; echo
  python bench.py --workload snapshot --no-cpu --no-e2e --steps 3 2>/dev/null | grep -o "\"kernel_ms\": [0-9.]*"
done
