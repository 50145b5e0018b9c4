#!/bin/bash
set -u
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_long.py -x -q 2>&1 | tail -5
GDY_PERSIST_PROF=1 GDY_PERSIST_NACT=2048 timeout 300 python tools/step_latency.py --n 1024 --steps 3000 2>&1 | head -5
GDY_PERSIST_PROF=1 GDY_PERSIST_NACT=2048 timeout 300 python tools/step_latency.py --n 16384 --steps 3000 2>&1
GDY_PERSIST_NACT=2048 timeout 300 python tools/step_latency.py --n 262144 --steps 1500 2>&1
