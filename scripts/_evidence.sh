# Round evidence on one B200: bench lines (both workloads + reference arm), ncu launch lists, ncu --set full of the two dominant kernels, all BASELINE configs.
set -x
python bench.py > gpurun_out/ev_bench_shell.json 2> gpurun_out/ev_bench_shell.err
python bench.py --workload snapshot > gpurun_out/ev_bench_snap.json 2> gpurun_out/ev_bench_snap.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/ev_bench_ref.json 2> gpurun_out/ev_bench_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/ev_launches_shell.csv python bench.py --no-cpu --no-e2e --steps 2 --warmup 1 > gpurun_out/ev_ncu1.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/ev_launches_snap.csv python bench.py --workload snapshot --no-cpu --no-e2e --steps 2 --warmup 1 > gpurun_out/ev_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:^gather_kernel -s 1 -c 1 -o gpurun_out/ev_gather python bench.py --no-cpu --no-e2e --steps 1 --warmup 1 > gpurun_out/ev_ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:snap_gather_kernel -s 1 -c 1 -o gpurun_out/ev_snap python bench.py --workload snapshot --no-cpu --no-e2e --steps 1 --warmup 1 > gpurun_out/ev_ncu4.log 2>&1
python scripts/measure_all.py c1,c2,c3,c5a,c5p,c5b,pk > gpurun_out/ev_measure_all.log 2>&1
tail -3 gpurun_out/ev_measure_all.log | cut -c1-400
