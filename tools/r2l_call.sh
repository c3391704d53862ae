#!/bin/bash
# r2l: FP64 peak, scale parity tests, phase breakdown of the chunked cell loop (ablations 1/2 timing only, 3 functional)
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 120 tools/bench_fp64 > $O/r2l_bench_fp64.txt 2>&1; echo "fp64 rc=$?"; cat $O/r2l_bench_fp64.txt
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv,noheader >> $O/r2l_bench_fp64.txt
timeout 600 python -m pytest tests/test_gpu_scale.py -m gpu -q > $O/r2l_pytest_scale.log 2>&1; echo "scale rc=$? $(tail -1 $O/r2l_pytest_scale.log)"
for t in default cabl1 cabl2 cabl3; do
  L=""; [ $t != default ] && L=$PWD/variants/libmgb200_$t.so
  for p in 4 3; do
    echo "variant $t MGB_CHUNKED=1 p=$p" >> $O/r2l_chunk_ablation.txt
    MGB_CHUNKED=1 MGB200_LIB=$L timeout 200 python tools/time_vmult.py --nref 7 --degree $p --reps 30 >> $O/r2l_chunk_ablation.txt 2>&1
  done
done
echo "atomic kernel (MGB_CHUNKED unset) p=4" >> $O/r2l_chunk_ablation.txt
timeout 200 python tools/time_vmult.py --nref 7 --degree 4 --reps 30 >> $O/r2l_chunk_ablation.txt 2>&1
grep -E "variant|atomic|cell_kernel_ms" $O/r2l_chunk_ablation.txt | sed 's/"n_dofs.*"vmult_ms"/"vmult_ms"/' | cut -c1-260
MGB_CHUNKED=1 MGB200_LIB=$PWD/variants/libmgb200_cabl3.so timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "vmult_and_diagonal or vcycle" > $O/r2l_pytest_cabl3.log 2>&1; echo "cabl3 parity rc=$? $(tail -1 $O/r2l_pytest_cabl3.log)"
