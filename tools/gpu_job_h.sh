#!/bin/bash
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -4
timeout 300 python tools/sweep_big.py > $O/sweep_big_v5.csv 2>&1; echo rc=$?
cat $O/sweep_big_v5.csv
timeout 300 python tools/step_latency.py --n 65536 --steps 3000 --spectrum salpeter > $O/step_latency_n64k_salpeter_v5b.txt 2>&1; echo rc=$?
head -6 $O/step_latency_n64k_salpeter_v5b.txt
