#!/bin/bash
# 8 GPUs: N=256k to t=1 with replicated tiny blocks (GDY_REPL_NACT default 16), then the bench line at 8 ranks
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
TR() { np=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $((29500 + np)) "$@"; }
timeout 400 bash -c "$(declare -f TR); TR 8 tools/integrate.py --particles 262144 --tend 1" > $O/integrate_n256k_8gpu_v10.json 2> $O/integrate_n256k_8gpu_v10.err; echo "integ rc=$?"
timeout 400 bash -c "$(declare -f TR); TR 8 bench.py --gpus 8 --steps 5 --warmup 3" > $O/bench_8gpu_v10.json 2> $O/bench_8gpu_v10.err; echo "bench rc=$?"
python - <<PY
import json
d = json.loads(open("$O/integrate_n256k_8gpu_v10.json").read().strip().splitlines()[-1])
print("wall", round(d["wall_s"], 3), "steps", d["block_steps"], "GF", round(d["gflops_60"]), "dE", d["cum_rel_energy_error"])
print(" ".join("%s:%.1f/%.2fs" % (b["nact"][1], b["median_us"], b["share_of_wall"] * d["wall_s"]) for b in d["bins"]))
d = json.loads(open("$O/bench_8gpu_v10.json").read().strip().splitlines()[-1])
print("bench8", round(d["value"]), d["ms_per_step"], d["e2e"]["value"], d["integration"]["us_per_block_step"])
PY
