#!/bin/bash
# 8 GPUs of one box: bench at 8/4/2 ranks, N = 1M full-active sweep at 8/1, N = 256k integration at 8 ranks
# (fused NVLink exchange vs NCCL), 2-GPU sharded tests
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
nvidia-smi topo -m > $O/topo_8gpu.txt 2>&1
TR() { np=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $((29500 + np)) "$@"; }
timeout 300 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q 2>&1 | tail -4
for g in 8 4 2; do
  timeout 300 bash -c "$(declare -f TR); TR $g bench.py --gpus $g --steps 5 --warmup 3" > $O/bench_${g}gpu_v4.json 2> $O/bench_${g}gpu_v4.err; echo bench$g rc=$?
done
timeout 200 python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_1gpu_v4b.json 2> $O/bench_1gpu_v4b.err; echo bench1 rc=$?
for g in 8 4 2; do
  timeout 300 bash -c "$(declare -f TR); TR $g bench.py --gpus $g --particles 1048576 --steps 3 --warmup 3 --no-integration" > $O/sweep1m_${g}gpu_v4.json 2> $O/sweep1m_${g}gpu_v4.err; echo sweep1m_$g rc=$?
done
timeout 300 bash -c "$(declare -f TR); TR 8 tools/integrate.py --particles 262144 --tend 1 --out $O/integ_256k_8gpu_peers_v4.json" > /dev/null 2> $O/integ_f.err; echo integ8 rc=$?
timeout 200 bash -c "$(declare -f TR); TR 8 tools/integrate.py --particles 262144 --tend 0.03125 --no-peers --out $O/integ_256k_8gpu_nccl_t32_v4.json" > /dev/null 2>> $O/integ_f.err; echo integ8nccl rc=$?
timeout 200 bash -c "$(declare -f TR); TR 8 tools/integrate.py --particles 262144 --tend 0.03125 --out $O/integ_256k_8gpu_peers_t32_v4.json" > /dev/null 2>> $O/integ_f.err; echo integ8peers rc=$?
timeout 200 bash -c "$(declare -f TR); TR 8 tools/integrate.py --particles 16384 --tend 1 --out $O/integ_16k_8gpu_peers_v4.json" > /dev/null 2>> $O/integ_f.err; echo integ16k8 rc=$?
tail -5 $O/integ_f.err
python - <<PY
import json,glob
for f in sorted(glob.glob("$O/bench_*gpu_v4*.json"))+sorted(glob.glob("$O/sweep1m_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f.split('/')[-1], d["n_gpus"], "GF", round(d["value"]), "ms", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["value"]), "integ", (d.get("integration") or {}).get("us_per_block_step"))
    except Exception as e: print(f, "ERR", e)
for f in sorted(glob.glob("$O/integ_*v4.json")):
    d=json.load(open(f)); print(f.split('/')[-1], "steps", d["block_steps"], "wall", round(d["wall_s"],3), "us/step", round(d["us_per_block_step"],1), "GF", round(d["gflops_60"]), "dE", d["cum_rel_energy_error"], "persist", d.get("persist_steps"))
PY
