#!/bin/bash
# new force plan (warp_i64 everywhere, ~24 waves of small CTAs): GPU tests, sizes sweep, bench, N=256k to t=1
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_v12.log 2>&1; echo pytest_rc=$?
tail -2 $O/pytest_gpu_v12.log
python tools/sweep_big.py --n 262144 --nacts 300,600,1172,2416,3636,4819,7000,10460,18652,40000,59491,131072,262144 --variants 0 --reps 5 | grep -E ",-1,-1," > $O/sweep_auto_v12.csv; cat $O/sweep_auto_v12.csv
timeout 600 python bench.py --no-cpu-baseline > $O/bench_1gpu_v12.json 2> $O/bench_1gpu_v12.err; echo bench_rc=$?
timeout 600 python tools/integrate.py --particles 262144 --tend 1 > $O/integrate_n256k_1gpu_v12.json 2> $O/integrate_n256k_1gpu_v12.err; echo integ_rc=$?
python - <<PY
import json
d=json.loads(open("$O/bench_1gpu_v12.json").read().strip().splitlines()[-1])
print("bench", round(d["value"]), round(d["roofline"]["frac"],4), round(d["e2e"]["value"]), d["integration"]["us_per_block_step"], d["gpu_launches"])
d=json.loads(open("$O/integrate_n256k_1gpu_v12.json").read().strip().splitlines()[-1])
print("integ wall", d["wall_s"], "steps", d["block_steps"], "gflops", d["gflops_60"], "dE", d["cum_rel_energy_error"])
print(" ".join("%s:%.0f/%.2fs" % (b["nact"][1], b["median_us"], b["share_of_wall"] * d["wall_s"]) for b in d["bins"]))
PY
