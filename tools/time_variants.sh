#!/bin/bash
# time_variants.sh <tag> ... [-- time_vmult.py options] -- on the GPU box: time the default library ("default") and
# the variants built by tools/build_variants.sh with tools/time_vmult.py (MGB200_LIB selects the library), then run the
# operator parity tests against every variant (NO_PYTEST=1 skips them).  Output: gpurun_out/variants_time.txt, gpurun_out/variants_pytest_<tag>.log
ROOT=$(cd "$(dirname "$0")/.." && pwd); cd "$ROOT"
O=gpurun_out; mkdir -p $O
tags=(); opts=()
while [ $# -gt 0 ]; do if [ "$1" = "--" ]; then shift; opts=("$@"); break; fi; tags+=("$1"); shift; done
[ ${#opts[@]} -eq 0 ] && opts=(--nref 7 --degree 4 --reps 50)
for t in default "${tags[@]}"; do
  L=""; [ $t != default ] && L=$ROOT/variants/libmgb200_$t.so
  echo "variant $t" >> $O/variants_time.txt
  MGB200_LIB=$L timeout 120 python tools/time_vmult.py "${opts[@]}" >> $O/variants_time.txt 2>&1
done
cat $O/variants_time.txt
[ -n "$NO_PYTEST" ] && exit 0   # timing-only variants (e.g. -DMGB_SLAB_ABLATE=n, wrong results by construction)
for t in "${tags[@]}"; do
  MGB200_LIB=$ROOT/variants/libmgb200_$t.so timeout 300 python -m pytest tests -m gpu -q -k "operator or vmult or general_geometry" > $O/variants_pytest_$t.log 2>&1
  echo "variant $t pytest rc=$? $(tail -1 $O/variants_pytest_$t.log)"
done
