#!/bin/bash
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
for n in 16384 65536 262144; do
  for cap in 256 512 1024 2048; do
    steps=3000; [ $n = 262144 ] && steps=1500
    echo "== N=$n cap=$cap"
    GDY_PERSIST_NACT=$cap timeout 300 python tools/step_latency.py --n $n --steps $steps 2>&1 | grep -E "^N=|\(   128|\(   256|\(   512|\(  1024"
  done
done
timeout 300 python bench.py --no-cpu-baseline --steps 3 > $O/bench_1gpu_v8.json 2> $O/bench_1gpu_v8.err; echo rc=$?
python -c "
import json;d=json.loads(open('$O/bench_1gpu_v8.json').read().strip().splitlines()[-1]);print(json.dumps(d['roofline_hbm']))"
