#!/bin/bash
# 8 GPUs: A/B of the replication threshold for tiny blocks, N=256k to t=1/32, per-step traces of rank 0
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
TR() { np=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $((29500 + np)) "$@"; }
for repl in 0 16 48; do
  GDY_REPL_NACT=$repl timeout 300 bash -c "$(declare -f TR); TR 8 tools/integrate.py --particles 262144 --tend 0.03125 --dump-trace $O/trace_8gpu_repl$repl.npz" > $O/trace_8gpu_repl$repl.json 2> $O/trace_8gpu_repl$repl.err; echo "repl $repl rc=$?"
done
python - <<PY
import json, numpy as np
for repl in (0, 16, 48):
    d = json.loads(open("$O/trace_8gpu_repl%d.json" % repl).read().strip().splitlines()[-1])
    t = np.load("$O/trace_8gpu_repl%d.npz" % repl)
    na, us = t["nact"], t["us"]
    print("repl", repl, "wall", round(d["wall_s"], 4), "steps", d["block_steps"], "small(<=512) ms", round(us[na <= 512].sum() / 1e3, 2),
          "<=16 ms", round(us[na <= 16].sum() / 1e3, 2), "17..512 ms", round(us[(na > 16) & (na <= 512)].sum() / 1e3, 2), "big ms", round(us[na > 512].sum() / 1e3, 2))
PY
