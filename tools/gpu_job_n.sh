#!/bin/bash
# ncu source-level capture of the full-active force kernel, production inner loop vs the software-pipelined form
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
for v in F0_U2_R512 F3_U1_R512; do
  GDY_LIB=$PWD/gravidy_b200/lab/libgdy_$v.so timeout 600 ncu --set full --clock-control none --import-source on \
     -k regex:k_force_warp -s 2 -c 1 -f -o $O/lab_$v python tools/sweep_big.py --n 262144 --nacts 262144 --variants 0 --reps 3 > $O/lab_ncu_$v.log 2>&1
  echo "$v rc=$?"
done
ls -la $O/*.ncu-rep
