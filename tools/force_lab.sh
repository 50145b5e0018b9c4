#!/bin/bash
# force_lab.sh build|run — inner-loop variants of k_force_warp side by side (GDY_FORCE_FORM : GDY_FORCE_UNROLL : GDY_FORCE_RESIDENT),
# each a full lab build of the library under gravidy_b200/lab/, timed with tools/sweep_big.py via GDY_LIB.
#   tools/force_lab.sh build "0:2 1:1 1:2:384"      (here: nvcc, prints the modelled FP64 cost from the SASS)
#   tools/force_lab.sh run   "0:2 1:1 1:2:384"      (on the GPU box)
set -u
cd "$(dirname "$0")/.."
mode=$1; cfgs=$2
for cfg in $cfgs; do
  IFS=: read f u r <<< "$cfg"; r=${r:-512}
  lib=gravidy_b200/lab/libgdy_F${f}_U${u}_R${r}.so
  if [ "$mode" = build ]; then
    ( cd gravidy_b200/csrc && nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC -Xcompiler -O2 \
        -DGDY_FORCE_FORM=$f -DGDY_FORCE_UNROLL=$u -DGDY_FORCE_RESIDENT=$r ${LAB_EXTRA:-} -o ../lab/$(basename $lib) gdy_ctx.cu -ldl ) || exit 1
    echo "form $f unroll $u resident $r: $(python tools/sass_fp64_cost.py $lib k_force_warpILi2ELi4ELi1ELi3ELb0 $((4*u)) | sed 's/.*loop of/loop of/')"
  else
    echo "== form $f unroll $u resident $r"
    GDY_LIB=$PWD/$lib timeout 300 python tools/sweep_big.py --n ${LAB_N:-262144} --nacts ${LAB_NACTS:-8192,262144} --variants 0 --reps 4 | grep -v ",0,1,"
  fi
done
