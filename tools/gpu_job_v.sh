#!/bin/bash
# BASELINE config 4 at G ranks (G = number of visible GPUs): N=256k to t=1 NBU, j-sharded; at G=1 also config 3 (N=64k Salpeter) and config 2 (N=16k)
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
G=$(nvidia-smi -L | wc -l)
TR() { np=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $((29500 + np)) "$@"; }
if [ $G -gt 1 ]; then
  timeout 600 bash -c "$(declare -f TR); TR $G tools/integrate.py --particles 262144 --tend 1" > $O/integrate_n256k_${G}gpu_v11.json 2> $O/integrate_n256k_${G}gpu_v11.err; echo "integ rc=$?"
  f=$O/integrate_n256k_${G}gpu_v11.json
else
  timeout 300 python tools/integrate.py --particles 65536 --spectrum salpeter --tend 1 > $O/integrate_n64k_salpeter_v11.json 2> $O/integrate_n64k_salpeter_v11.err; echo "n64k rc=$?"
  timeout 300 python tools/integrate.py --particles 16384 --tend 1 > $O/integrate_n16k_v11.json 2> $O/integrate_n16k_v11.err; echo "n16k rc=$?"
  f="$O/integrate_n64k_salpeter_v11.json $O/integrate_n16k_v11.json"
fi
python - $f <<PY
import json, sys
for f in sys.argv[1:]:
    d = json.loads(open(f).read().strip().splitlines()[-1])
    print(f, "n", d["n"], "gpus", d["n_gpus"], "wall", round(d["wall_s"], 3), "steps", d["block_steps"], "GF", round(d["gflops_60"]), "dE", d["cum_rel_energy_error"], "persist", d.get("persist_steps"))
PY
