#!/bin/bash
# r2g: chunk-size / residency variants of the chunked cell loop at L=7 p=4
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
for t in default minb2 nc16b7 nc16b5; do
  L=""; [ $t != default ] && L=$PWD/variants/libmgb200_$t.so
  echo "variant $t" >> $O/r2g_variants.txt
  MGB200_LIB=$L timeout 200 python tools/time_vmult.py --nref 7 --degree 4 --reps 30 >> $O/r2g_variants.txt 2>&1
  MGB200_LIB=$L timeout 300 python -m pytest tests -m gpu -q -x -k "vmult_and_diagonal" 2>&1 | tail -1 >> $O/r2g_variants.txt
done
cat $O/r2g_variants.txt
