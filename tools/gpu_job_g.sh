#!/bin/bash
# 1 GPU: multi-warp candidate scan of the persistent kernel (v5): tests, per-step latency, ncu captures
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
GDY_PERSIST_PROF=1 timeout 300 python tools/step_latency.py --n 262144 --steps 1500 > $O/step_latency_n256k_v5_prof.txt 2>&1; echo rc=$?
timeout 300 python tools/step_latency.py --n 262144 --steps 1500 > $O/step_latency_n256k_v5.txt 2>&1; echo rc=$?
timeout 300 python tools/step_latency.py --n 65536 --steps 3000 --spectrum salpeter > $O/step_latency_n64k_salpeter_v5.txt 2>&1; echo rc=$?
timeout 300 python tools/step_latency.py --n 16384 --steps 3000 > $O/step_latency_n16k_v5.txt 2>&1; echo rc=$?
timeout 300 python tools/step_latency.py --n 1048576 --steps 600 > $O/step_latency_n1m_v5.txt 2>&1; echo rc=$?
head -22 $O/step_latency_n256k_v5_prof.txt
head -12 $O/step_latency_n64k_salpeter_v5.txt
head -3 $O/step_latency_n1m_v5.txt
# ncu: the DFMA peak micro-benchmark (is 91.7 % of nominal the pipe's ceiling?) and one persistent launch
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_dfma_peak -s 40 -c 2 -o $O/prof_dfma_peak python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-integration > $O/ncu_dfma.log 2>&1; echo rc=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_persist_steps -s 20 -c 1 -o $O/prof_persist python tools/step_latency.py --n 262144 --steps 600 > $O/ncu_persist.log 2>&1; echo rc=$?
ls -la $O/*.ncu-rep
