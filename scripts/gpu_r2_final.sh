#!/bin/bash
# round 2, final call: the round's evidence on one B200 -- GPU tests, the four bench workloads + reference arm, ncu launch
# lists of the bench commands, ncu --set full of the dominant kernels (each only after its plain command exited 0)
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > $O/r2f_smi.txt; nproc >> $O/r2f_smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q --durations=10 > $O/r2f_pytest.log 2>&1; echo "pytest rc=$?"; tail -16 $O/r2f_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/r2f_bench_shell_n1.json 2> $O/r2f_bench_shell_n1.err; echo "bench shell rc=$?"
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > $O/r2f_bench_reference_arm.json 2> $O/r2f_bench_reference_arm.err; echo "ref arm rc=$?"
for w in paint snapshot; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 > $O/r2f_bench_${w}_n1.json 2> $O/r2f_bench_${w}_n1.err; echo "bench $w rc=$?"
done
timeout 900 python bench.py --workload grid3d --steps 3 --warmup 2 > $O/r2f_bench_grid3d_n1.json 2> $O/r2f_bench_grid3d_n1.err; echo "bench grid3d rc=$?"
# launch lists (per-launch durations of the same commands, serialised by ncu: shares, not absolutes)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2f_launches_bench_shell.csv \
  python bench.py --no-cpu --no-e2e --steps 2 --warmup 1 > $O/r2f_ncu1.log 2>&1; echo "launch list shell rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2f_launches_bench_snapshot.csv \
  python bench.py --workload snapshot --no-cpu --no-e2e --steps 2 --warmup 1 > $O/r2f_ncu2.log 2>&1; echo "launch list snapshot rc=$?"
# --set full captures
timeout 600 ncu --set full --clock-control none --import-source on -k regex:^gather_kernel -s 1 -c 1 -o $O/r2f_gather \
  python bench.py --no-cpu --no-e2e --steps 1 --warmup 1 > $O/r2f_ncu3.log 2>&1; echo "ncu gather rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:snap_stream_kernel -s 1 -c 1 -o $O/r2f_snap \
  python bench.py --workload snapshot --no-cpu --no-e2e --steps 1 --warmup 1 > $O/r2f_ncu4.log 2>&1; echo "ncu snap rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:^rg_ -s 2 -c 2 -o $O/r2f_regrid \
  python bench.py --no-cpu --no-e2e --steps 1 --warmup 1 > $O/r2f_ncu5.log 2>&1; echo "ncu regrid rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:^gather_kernel -s 1 -c 1 -o $O/r2f_gather_paint \
  python bench.py --workload paint --no-cpu --no-e2e --steps 1 --warmup 1 > $O/r2f_ncu6.log 2>&1; echo "ncu paint rc=$?"
python - <<'PY'
import json
for w in ("shell", "paint", "snapshot", "grid3d"):
    try:
        d = json.load(open(f"gpurun_out/r2f_bench_{w}_n1.json"))
        print(w, round(d["ms_per_step"], 2), "ms/step", "%.3g" % d["value"], d["unit"], "roofline", round(d["roofline"]["frac"], 3),
              "e2e ms", d["e2e"] and round(d["e2e"]["ms"], 1), d["config"].get("stages_ms_max_over_ranks"))
    except Exception as e:
        print(w, "failed", e)
PY
