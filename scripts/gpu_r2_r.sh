#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
timeout 100 $TR scripts/check_sharded.py 512 20000 > gpurun_out/r2r_check_n2.log 2>&1; echo "check rc=$?"; grep -E "world|SHARDED|rror" gpurun_out/r2r_check_n2.log | tail -12
