#!/bin/bash
# persistent small-step kernel: parity, latency, phase profile
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
echo "== pytest fused/long"; timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_long.py -x -q > $O/pytest_b.log 2>&1; echo "rc=$?"; tail -15 $O/pytest_b.log
for nm in 0 64 512 2048; do
  echo "== step latency persist=$nm"
  GDY_PERSIST_NACT=$nm timeout 300 python tools/step_latency.py --n 262144 --steps 1500 > $O/lat_256k_p$nm.txt 2>&1; head -4 $O/lat_256k_p$nm.txt
  GDY_PERSIST_NACT=$nm timeout 300 python tools/step_latency.py --n 16384 --steps 3000 > $O/lat_16k_p$nm.txt 2>&1; head -4 $O/lat_16k_p$nm.txt
done
echo "== phase profile"
GDY_PERSIST_PROF=1 timeout 300 python tools/step_latency.py --n 16384 --steps 3000 2>&1 | head -4
GDY_PERSIST_PROF=1 timeout 300 python tools/step_latency.py --n 262144 --steps 1500 2>&1 | head -4
GDY_PERSIST_PROF=1 GDY_PERSIST_NACT=8 timeout 300 python tools/step_latency.py --n 16384 --steps 3000 2>&1 | head -4
timeout 300 python tools/integrate.py --n 16384 --t 1 --out $O/integrate_n16k_p.json > /dev/null 2> $O/integrate_b.err; echo "rc=$?"
python -c "
import json; d=json.load(open('$O/integrate_n16k_p.json')); print('n16k', d['block_steps'], d['sum_nact'], d['wall_s'], d['gflops_60'], d['cum_rel_energy_error'])"
