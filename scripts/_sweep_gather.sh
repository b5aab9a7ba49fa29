# SWEEP="flags1|flags2|..." ; kernel_ms of the shell gather kernel for a few build variants (run on the GPU box)
IFS="|"; for v in ${SWEEP:-""}; do IFS=" "
  echo "== [$v]"; BF_NVCC_EXTRA="$v" python -m baryonification_b200.build > /tmp/b.log 2>&1; grep -A3 "gather_kernelILi1EEv11gather_args" baryonification_b200/build.log | grep -E "Used" | tr '\n' ' '; echo
  BF_NVCC_EXTRA="$v" python bench.py --no-cpu --no-e2e --steps 3 2>/dev/null | grep -o "\"kernel_ms\": [0-9.]*\|\"ms_per_step\": [0-9.]*" | tr '\n' ' '; echo
done
