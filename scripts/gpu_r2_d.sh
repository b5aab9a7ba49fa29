#!/bin/bash
# round 2, call D (2 GPUs): sharded runners vs single GPU on NCCL, bench of the workloads at N=2
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
timeout 900 $TR scripts/check_sharded.py 1024 50000 > gpurun_out/r2d_check.log 2>&1; echo "check rc=$?"; grep -E "world|SHARDED|Error|error" gpurun_out/r2d_check.log | tail -20
for w in shell paint snapshot; do
  timeout 900 $TR bench.py --gpus 2 --workload $w --steps 10 --warmup 3 > gpurun_out/r2d_bench_${w}_n2.json 2> gpurun_out/r2d_bench_${w}_n2.err; echo "bench $w rc=$?"
  tail -c 2500 gpurun_out/r2d_bench_${w}_n2.json; grep -v "^W\|^\[W\|NCCL version" gpurun_out/r2d_bench_${w}_n2.err | tail -5
done
timeout 900 $TR bench.py --gpus 2 --workload grid3d --ngrid 512 --nhalo 20000 --steps 3 --warmup 2 > gpurun_out/r2d_bench_grid3d512_n2.json 2> gpurun_out/r2d_bench_grid3d512_n2.err; echo "grid3d rc=$?"
tail -c 2500 gpurun_out/r2d_bench_grid3d512_n2.json; tail -5 gpurun_out/r2d_bench_grid3d512_n2.err
