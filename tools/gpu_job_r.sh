#!/bin/bash
# per-step traces: 1 GPU vs 2 GPUs sharded (GDY_REPL_NACT 0 / 16), N=256k to t=1/64
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
TR() { np=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $((29500 + np)) "$@"; }
timeout 300 python tools/integrate.py --particles 262144 --tend 0.015625 --dump-trace $O/trace_1gpu.npz > $O/trace_1gpu.json 2> $O/trace_1gpu.err; echo rc=$?
for repl in 0 16; do
  GDY_REPL_NACT=$repl timeout 300 bash -c "$(declare -f TR); TR 2 tools/integrate.py --particles 262144 --tend 0.015625 --dump-trace $O/trace_2gpu_repl$repl.npz" > $O/trace_2gpu_repl$repl.json 2> $O/trace_2gpu_repl$repl.err; echo "repl $repl rc=$?"
done
