#!/bin/bash
# persistent small-step kernel: parity, sanitizer, latency
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
echo "== pytest fused/long"; timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_long.py -x -q > $O/pytest_b.log 2>&1; echo "rc=$?"; tail -15 $O/pytest_b.log
echo "== sanitizer memcheck"; timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_parity.py -x -q -k "fused_run and n97" > $O/sanitizer_mem.log 2>&1; echo "rc=$?"; tail -5 $O/sanitizer_mem.log
echo "== sanitizer racecheck"; timeout 600 compute-sanitizer --tool racecheck --error-exitcode 9 python -m pytest tests/test_gpu_parity.py -x -q -k "fused_run and n97 and 512" > $O/sanitizer_race.log 2>&1; echo "rc=$?"; tail -5 $O/sanitizer_race.log
for nm in 0 8 64 256 512; do
  echo "== step latency persist=$nm"
  GDY_PERSIST_NACT=$nm timeout 300 python tools/step_latency.py --n 262144 --steps 1500 > $O/lat_256k_p$nm.txt 2>&1; cat $O/lat_256k_p$nm.txt
  GDY_PERSIST_NACT=$nm timeout 300 python tools/step_latency.py --n 16384 --steps 3000 > $O/lat_16k_p$nm.txt 2>&1; cat $O/lat_16k_p$nm.txt
done
timeout 300 python tools/integrate.py --n 16384 --t 1 --out $O/integrate_n16k_p.json > /dev/null 2> $O/integrate_b.err; echo "rc=$?"
python -c "
import json; d=json.load(open('$O/integrate_n16k_p.json')); print('n16k', d['block_steps'], d['sum_nact'], d['wall_s'], d['gflops_60'], d['cum_rel_energy_error'])"
