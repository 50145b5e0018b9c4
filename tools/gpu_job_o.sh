#!/bin/bash
# after the software-pipelined inner loop (GDY_FORCE_FORM 3): GPU tests, bench, reuse micro-benchmark, N=256k to t=1
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_v10.log 2>&1; echo pytest_rc=$?
tail -3 $O/pytest_gpu_v10.log
tools/fp64_reuse > $O/fp64_reuse_v9.txt 2>&1; cat $O/fp64_reuse_v9.txt
timeout 600 python bench.py > $O/bench_1gpu_v10.json 2> $O/bench_1gpu_v10.err; echo bench_rc=$?
timeout 600 python tools/integrate.py --particles 262144 --tend 1 > $O/integrate_n256k_1gpu_v10.json 2> $O/integrate_n256k_1gpu_v10.err; echo integ_rc=$?
python - <<PY
import json
d=json.loads(open("$O/bench_1gpu_v10.json").read().strip().splitlines()[-1])
print("bench", round(d["value"]), d["roofline"]["frac"], d["e2e"]["value"], d["integration"]["us_per_block_step"])
d=json.loads(open("$O/integrate_n256k_1gpu_v10.json").read().strip().splitlines()[-1])
print("integ wall", d["wall_s"], "steps", d["block_steps"], "gflops", d["gflops_60"], "dE", d["cum_rel_energy_error"])
PY
