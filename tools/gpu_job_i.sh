#!/bin/bash
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python tools/sweep_big.py --variants 0 > $O/sweep_big_v6.csv 2>&1; echo rc=$?
cat $O/sweep_big_v6.csv
timeout 600 python bench.py --no-cpu-baseline > $O/bench_1gpu_v6.json 2> $O/bench_1gpu_v6.err; echo rc=$?
python -c "
import json;d=json.loads(open('$O/bench_1gpu_v6.json').read().strip().splitlines()[-1]);print(d['value'],d['roofline'],d['integration']['us_per_block_step'])"
