#!/bin/bash
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
timeout 300 python -m pytest tests/test_gpu_legacy.py -m gpu -x -q 2>&1 | tail -5
timeout 300 python tools/legacy_gpu.py --n 16384 --nacts 32,256,2048 --out $O/legacy_gpu_n16k.json
timeout 600 python tools/legacy_gpu.py --n 262144 --nacts 32,512,8192,65536 --reps 2 --out $O/legacy_gpu_n256k.json
