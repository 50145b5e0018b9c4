#!/bin/bash
# 2-GPU: sharded tests (NCCL + peer path), integration with and without the fused exchange
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
nvidia-smi topo -m > $O/topo.txt 2>&1
echo skip-tests
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tools/integrate.py --particles 16384 --tend 0.25 --out $O/integ_16k_2gpu_peers.json > /dev/null 2> $O/integ_d.err; echo rc=$?
timeout 600 $TR tools/integrate.py --particles 16384 --tend 0.25 --no-peers --out $O/integ_16k_2gpu_nccl.json > /dev/null 2>> $O/integ_d.err; echo rc=$?
timeout 600 python tools/integrate.py --particles 16384 --tend 0.25 --out $O/integ_16k_1gpu.json > /dev/null 2>> $O/integ_d.err; echo rc=$?
timeout 600 $TR tools/integrate.py --particles 262144 --tend 0.015625 --out $O/integ_256k_2gpu_peers.json > /dev/null 2>> $O/integ_d.err; echo rc=$?
timeout 600 $TR tools/integrate.py --particles 262144 --tend 0.015625 --no-peers --out $O/integ_256k_2gpu_nccl.json > /dev/null 2>> $O/integ_d.err; echo rc=$?
timeout 600 python tools/integrate.py --particles 262144 --tend 0.015625 --out $O/integ_256k_1gpu.json > /dev/null 2>> $O/integ_d.err; echo rc=$?
tail -5 $O/integ_d.err
python - <<PY
import json,glob
for f in sorted(glob.glob("$O/integ_*.json")):
    d=json.load(open(f)); print(f.split('/')[-1], "steps", d["block_steps"], "wall", round(d["wall_s"],3), "us/step", round(d["us_per_block_step"],1), "GF", round(d["gflops_60"]), "dE", d["cum_rel_energy_error"], "persist", d.get("persist_steps"))
PY
