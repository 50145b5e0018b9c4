#!/bin/bash
# 1 GPU: full gpu test suite, bench (both arms), per-step latency with the persistent kernel, launch list
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref_v4.json 2> $O/bench_ref_v4.err; echo rc=$?
timeout 600 python bench.py > $O/bench_1gpu_v4.json 2> $O/bench_1gpu_v4.err; echo rc=$?
tail -c 1500 $O/bench_1gpu_v4.json
timeout 300 python tools/step_latency.py --n 16384 --steps 3000 > $O/step_latency_n16k_v4.txt 2>&1; echo rc=$?
GDY_PERSIST_NACT=0 timeout 300 python tools/step_latency.py --n 16384 --steps 3000 > $O/step_latency_n16k_v4_nopersist.txt 2>&1; echo rc=$?
timeout 300 python tools/step_latency.py --n 262144 --steps 1500 > $O/step_latency_n256k_v4.txt 2>&1; echo rc=$?
timeout 300 python tools/step_latency.py --n 65536 --steps 3000 --spectrum salpeter > $O/step_latency_n64k_salpeter_v4.txt 2>&1; echo rc=$?
head -20 $O/step_latency_n16k_v4.txt
timeout 600 python tools/integrate.py --particles 16384 --tend 1 --out $O/integ_16k_1gpu_v4.json > /dev/null 2> $O/integ_e.err; echo rc=$?
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_bench_v4.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --integration-steps 100 > $O/ncu_bench.log 2>&1; echo rc=$?
python - <<PY
import json,glob
for f in sorted(glob.glob("$O/integ_*v4.json")):
    d=json.load(open(f)); print(f.split('/')[-1], "steps", d["block_steps"], "wall", round(d["wall_s"],3), "us/step", round(d["us_per_block_step"],1), "GF", round(d["gflops_60"]), "dE", d["cum_rel_energy_error"], "persist", d.get("persist_steps"))
PY
