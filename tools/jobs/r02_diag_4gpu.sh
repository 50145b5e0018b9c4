#!/bin/bash
# 4 GPUs: why do replicated small steps get slower with the GPU count?  (a) default sharded run with the phase profile of
# every rank, (b) everything replicated (no exchange at all, peers mapped), (c) four independent one-GPU runs side by side
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
export GDY_PERSIST_PROF=1
T=0.0078125
TR() { np=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $((29500 + np)) "$@"; }
timeout 300 python tools/integrate.py --particles 262144 --tend $T --rank-prof $O/diag_1gpu > $O/diag_1gpu.json 2> $O/diag_1gpu.err; echo "1gpu rc=$?"
timeout 300 bash -c "$(declare -f TR); TR 4 tools/integrate.py --particles 262144 --tend $T --rank-prof $O/diag_4gpu_default" > $O/diag_4gpu_default.json 2> $O/diag_4gpu_default.err; echo "default rc=$?"
GDY_REPL_NACT=2048 timeout 300 bash -c "$(declare -f TR); TR 4 tools/integrate.py --particles 262144 --tend $T --rank-prof $O/diag_4gpu_repl" > $O/diag_4gpu_repl.json 2> $O/diag_4gpu_repl.err; echo "repl rc=$?"
for g in 0 1 2 3; do
  CUDA_VISIBLE_DEVICES=$g timeout 300 python tools/integrate.py --particles 262144 --tend $T --rank-prof $O/diag_indep$g > $O/diag_indep$g.json 2> $O/diag_indep$g.err &
done
wait; echo "indep done"
nvidia-smi topo -m > $O/topo.txt 2>&1
tail -n 30 $O/diag_*.txt
