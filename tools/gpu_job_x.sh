#!/bin/bash
# ncu evidence for the v12 plan: full capture of the sweep kernel inside bench.py, then the launch list of the bench command
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:k_force_warp -s 4 -c 1 -f -o $O/prof_force_i64_v12 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-integration > $O/ncu_full_v12.log 2>&1; echo "full rc=$?"
timeout 700 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_bench_v12.csv python bench.py --no-cpu-baseline > $O/ncu_bench_v12.log 2>&1; echo "list rc=$?"
ls -la $O/prof_force_i64_v12.ncu-rep $O/launches_bench_v12.csv
