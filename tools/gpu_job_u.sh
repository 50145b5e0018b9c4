#!/bin/bash
# round-end validation on one GPU: GPU tests, all-kernel smoke, both bench arms, ncu launch list of the bench command
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_v11.log 2>&1; echo pytest_rc=$?
tail -2 $O/pytest_gpu_v11.log
timeout 200 python tools/sanitize_smoke.py 2>&1 | tail -1
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference_v11.json 2> $O/bench_reference_v11.err; echo ref_rc=$?
timeout 600 python bench.py > $O/bench_1gpu_v11.json 2> $O/bench_1gpu_v11.err; echo bench_rc=$?
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_bench_v11.csv python bench.py --no-cpu-baseline > $O/ncu_bench_v11.log 2>&1; echo ncu_rc=$?
python - <<PY
import json
d=json.loads(open("$O/bench_1gpu_v11.json").read().strip().splitlines()[-1])
print("bench", round(d["value"]), round(d["roofline"]["frac"],4), round(d["e2e"]["value"]), d["roofline_hbm"]["bandwidth_regime"], d["cpu_baseline"]["value"], d["cpu_baseline"]["loop"]["gpu_over_cpu_wall"], d["gpu_launches"], d["clocks"])
d=json.loads(open("$O/bench_reference_v11.json").read().strip().splitlines()[-1])
print("ref", d["value"], d["cpu_baseline"]["cores"])
PY
