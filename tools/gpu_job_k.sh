#!/bin/bash
# 8 GPUs: the north star's full-active N = 1M sweep at 8 and 1 ranks
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
TR() { np=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $((29500 + np)) "$@"; }
timeout 300 bash -c "$(declare -f TR); TR 8 bench.py --gpus 8 --particles 1048576 --steps 3 --warmup 3 --no-integration" > $O/sweep1m_8gpu_v6.json 2> $O/sweep1m_8gpu_v6.err; echo rc=$?
timeout 300 python bench.py --gpus 1 --particles 1048576 --steps 3 --warmup 3 --no-integration --no-cpu-baseline > $O/sweep1m_1gpu_v6.json 2> $O/sweep1m_1gpu_v6.err; echo rc=$?
tail -2 $O/sweep1m_8gpu_v6.err
python - <<PY
import json
for f in ["$O/sweep1m_8gpu_v6.json","$O/sweep1m_1gpu_v6.json"]:
    d=json.loads(open(f).read().strip().splitlines()[-1]); print(d["n_gpus"], d["config"]["n"], "GF", round(d["value"]), "ms", round(d["ms_per_step"],2), "frac", round(d["frac_fp64_peak"],3), "e2e", round(d["e2e"]["value"]))
PY
