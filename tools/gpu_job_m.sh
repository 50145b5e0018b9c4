#!/bin/bash
# validation after a container restore: GPU tests, both bench arms, N=256k to t=1 on one GPU with the persistent kernel
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_v9.log 2>&1; echo pytest_rc=$?
tail -3 $O/pytest_gpu_v9.log
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference_v9.json 2> $O/bench_reference_v9.err; echo ref_rc=$?
timeout 600 python bench.py > $O/bench_1gpu_v9.json 2> $O/bench_1gpu_v9.err; echo bench_rc=$?
tail -c 600 $O/bench_1gpu_v9.err
timeout 600 python tools/integrate.py --particles 262144 --tend 1 > $O/integrate_n256k_1gpu_v9.json 2> $O/integrate_n256k_1gpu_v9.err; echo integ_rc=$?
python - <<PY
import json
d=json.loads(open("$O/bench_1gpu_v9.json").read().strip().splitlines()[-1])
print("bench", round(d["value"]), d["roofline"]["frac"], d["e2e"]["value"], d.get("cpu_baseline"))
d=json.loads(open("$O/integrate_n256k_1gpu_v9.json").read().strip().splitlines()[-1])
print("integ wall", d["wall_s"], "steps", d["block_steps"], "gflops", d["gflops_60"], "dE", d["cum_rel_energy_error"])
PY
