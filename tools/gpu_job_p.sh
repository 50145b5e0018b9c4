#!/bin/bash
# 2 GPUs: sharded tests with replicated tiny blocks, then N=256k to t=1/64 with the replication threshold swept
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_host_program.py -m gpu -x -q > $O/pytest_2gpu_v10.log 2>&1; echo pytest_rc=$?
tail -3 $O/pytest_2gpu_v10.log
TR() { np=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $((29500 + np)) "$@"; }
for repl in 0 4 16 64; do
  GDY_REPL_NACT=$repl timeout 300 bash -c "$(declare -f TR); TR 2 tools/integrate.py --particles 262144 --tend 0.015625" > $O/integ_2gpu_repl$repl.json 2> $O/integ_2gpu_repl$repl.err; echo "repl $repl rc=$?"
done
python - <<PY
import json
for repl in (0, 4, 16, 64):
    d = json.loads(open("$O/integ_2gpu_repl%d.json" % repl).read().strip().splitlines()[-1])
    print("repl", repl, "wall", round(d["wall_s"], 4), "steps", d["block_steps"], "E", d["e1"],
          " ".join("%s:%.1f" % (b["nact"][1], b["median_us"]) for b in d["bins"] if b["nact"][1] <= 128 and b["steps"]))
PY
