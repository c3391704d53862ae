#!/bin/bash
# build_variants.sh "tag:flags" ... -- build kernel variants of libmgb200.so HERE (nvcc cross-compiles, in parallel), one per
# argument, into variants/libmgb200_<tag>.so (git-ignored, travels to the GPU box with gpurun).  flags are extra nvcc
# options, typically the -D switches of the cell kernels:
#   -DMGB_SLAB_W=<warps per block> -DMGB_SLAB_MINB=<blocks per SM> -DMGB_SLAB_LOCKSTEP -DMGB_SLAB_ABLATE=1|2|3   (operator.cuh)
#   -DMGB_Q1_W= -DMGB_Q1_MINB= -DMGB_Q1_NO_LOCKSTEP -DMGB_Q1_REG_METRIC -DMGB_Q1_EXACT_DIV -DMGB_Q1_FLOAT_METRIC   (operator_q1.cuh)
# Then, in ONE gpurun call:  tools/time_variants.sh <tag> ... [-- time_vmult.py options]
# example: tools/build_variants.sh "lock:-DMGB_SLAB_LOCKSTEP" "w8:-DMGB_SLAB_W=8 -DMGB_SLAB_MINB=1"
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
mkdir -p "$ROOT/variants"
for v in "$@"; do
  tag=${v%%:*}; fl=${v#*:}
  ( d=$(mktemp -d /tmp/mgb_var_${tag}_XXXX)
    make -C "$ROOT/dealii_multigrid_b200/csrc" BUILD=$d/build LIBDIR=$d/lib EXTRA="$fl" $d/lib/libmgb200.so > $d/log 2>&1 \
      && cp $d/lib/libmgb200.so "$ROOT/variants/libmgb200_$tag.so" && echo "built $tag ($fl)" || { echo "FAILED $tag: see $d/log"; tail -5 $d/log; } ) &
done
wait
ls -la "$ROOT/variants"
