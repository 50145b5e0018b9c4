#!/bin/bash
# compute-sanitizer memcheck over every kernel family (small N), then the round's final validation on one GPU
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
timeout 200 python tools/sanitize_smoke.py > $O/sanitize_plain.log 2>&1; echo "plain rc=$?"; tail -2 $O/sanitize_plain.log
timeout 420 compute-sanitizer --tool memcheck --print-limit 20 python tools/sanitize_smoke.py > $O/sanitize_memcheck_v10.log 2>&1; echo "memcheck rc=$?"
grep -E "ERROR SUMMARY|sanitize_smoke ok|Invalid|Error" $O/sanitize_memcheck_v10.log | head -12
