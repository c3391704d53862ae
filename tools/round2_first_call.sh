#!/bin/bash
# round2_first_call.sh -- the measurements queued at the end of round 1 (no GPU minutes were left), to be run as the FIRST
# gpurun call of the next round (about 7 minutes of box time):
#   here, before the call:   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/bench_bulk tools/bench_bulk.cu
#                            nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/bench_bulk_cells tools/bench_bulk_cells.cu
#                            python tools/span_table.py --nref 6        (writes tools/spans_L6_p4.bin, 63 MB)
#                            tools/build_variants.sh "lock:-DMGB_SLAB_LOCKSTEP" "q1minb3:-DMGB_Q1_MINB=3" "q1nolock:-DMGB_Q1_NO_LOCKSTEP" "q1fm:-DMGB_Q1_FLOAT_METRIC" \
#                                "abl1:-DMGB_SLAB_ABLATE=1" "abl2:-DMGB_SLAB_ABLATE=2" "abl3:-DMGB_SLAB_ABLATE=3"
#   gpurun --timeout 700 -- tools/round2_first_call.sh
# 1. short bulk copies / f64 bulk reduce-add per SM (does the TMA path carry a cell's ~10+10 spans?)
# 2. block-lockstep variant of the Cartesian vmult kernel against the default (L=7 p=4), with the operator parity tests
# 3. general-geometry kernel: MIN_BLOCKS=3 and no-lockstep variants at p=3 and p=4, and an ncu capture of the default
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 150 tools/bench_bulk 2000 > $O/r2a_bench_bulk.txt 2>&1; echo "bench_bulk rc=$?"
# 1b. the same on the real spans of the L=6 p=4 mesh: one pass over all cells, gather and gather+scatter
for w in 4 2; do timeout 120 tools/bench_bulk_cells tools/spans_L6_p4.bin $w >> $O/r2a_bench_bulk_cells.txt 2>&1; done; echo "bench_bulk_cells rc=$?"
tools/time_variants.sh lock -- --nref 7 --degree 4 --reps 50 > $O/r2a_variants_vmult.log 2>&1; mv $O/variants_time.txt $O/r2a_variants_vmult_time.txt
# 2b. where the cell kernel's time goes: no scatter / no gather / gather+scatter only (timing-only variants)
NO_PYTEST=1 tools/time_variants.sh abl1 abl2 abl3 -- --nref 7 --degree 4 --reps 50 > $O/r2a_ablation.log 2>&1; mv $O/variants_time.txt $O/r2a_ablation_time.txt
for p in 4 3; do
  tools/time_variants.sh q1minb3 q1nolock -- --nref 6 --degree $p --deform smooth --reps 20 > $O/r2a_variants_q1_p$p.log 2>&1
  mv $O/variants_time.txt $O/r2a_variants_q1_p${p}_time.txt
done
# 2c. locally refined mesh: cells with hanging-node constraints in a launch of their own (MGB_SPLIT_HANGING=1)
for v in 0 1; do
  echo "MGB_SPLIT_HANGING=$v" >> $O/r2a_split_hanging.txt
  MGB_SPLIT_HANGING=$v timeout 120 python tools/time_vmult.py --geometry quadrant --nref 7 --degree 4 --reps 50 >> $O/r2a_split_hanging.txt 2>&1
done
MGB_SPLIT_HANGING=1 timeout 300 python -m pytest tests -m gpu -q -x > $O/r2a_split_hanging_pytest.log 2>&1; echo "split-hanging pytest rc=$? $(tail -1 $O/r2a_split_hanging_pytest.log)"
# 3b. float kernels with the metric in FP32 (parity tolerance of the float test: 2e-5)
tools/time_variants.sh q1fm -- --nref 6 --degree 4 --number float --deform smooth --reps 20 > $O/r2a_variants_q1_float.log 2>&1
mv $O/variants_time.txt $O/r2a_variants_q1_float_time.txt
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_laplace_q1_slab --launch-skip 3 -c 1 -f -o $O/r2a_q1_slab_L6 \
  python tools/time_vmult.py --nref 6 --degree 4 --deform smooth --reps 2 > $O/r2a_ncu_q1.log 2>&1; echo "ncu rc=$?"
tail -n 30 $O/r2a_bench_bulk.txt $O/r2a_bench_bulk_cells.txt $O/r2a_variants_vmult_time.txt $O/r2a_ablation_time.txt $O/r2a_variants_q1_p4_time.txt
