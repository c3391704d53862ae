#!/bin/bash
# gpu_round.sh <tag> -- the per-change GPU pass: parity tests, smoke, bench lines, ncu launch list, one full ncu capture.
# Run under gpurun from the repo root; everything it writes goes to gpurun_out/<tag>_*.
TAG=${1:-run}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest.log
timeout 300 python __graft_entry__.py smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/${TAG}_smoke.log
for spec in "hypercube 6 4 double" "hypercube 7 4 double" "hypercube 7 4 float" "hypercube 7 1 double" "hypercube 7 2 double" "hypercube 7 3 double" "quadrant 7 4 double"; do
  set -- $spec
  timeout 300 python tools/time_vmult.py --geometry $1 --nref $2 --degree $3 --number $4 --reps 100 >> $O/${TAG}_time_vmult.txt 2>&1
done
# general-geometry kernel (mgb_operator_set_cell_vertices): smoothly deformed cells
for spec in "6 4 double" "6 3 double" "6 2 double" "6 1 double" "6 4 float" "7 4 double"; do
  set -- $spec
  timeout 300 python tools/time_vmult.py --nref $1 --degree $2 --number $3 --deform smooth --reps 20 >> $O/${TAG}_time_vmult_q1.txt 2>&1
done
timeout 600 python bench.py --nref 6 > $O/${TAG}_bench_L6.json 2> $O/${TAG}_bench_L6.err; echo "bench L6 rc=$?"
timeout 900 python bench.py > $O/${TAG}_bench_L7.json 2> $O/${TAG}_bench_L7.err; echo "bench L7 rc=$?"
timeout 600 python bench.py --number float --no-cpu > $O/${TAG}_bench_L7_float.json 2> $O/${TAG}_bench_L7_float.err; echo "bench L7 float rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_ref.json 2> $O/${TAG}_bench_ref.err; echo "bench ref rc=$?"
tail -c 600 $O/${TAG}_bench_L7.json
# ncu passes: never a bench value
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $O/${TAG}_launches_L6.csv \
  python bench.py --nref 6 --steps 1 --warmup 3 --no-cpu > $O/${TAG}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_laplace_apply --launch-skip 3 -c 1 -f -o $O/${TAG}_vmult_L6 \
  python tools/time_vmult.py --nref 6 --degree 4 --reps 2 > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
# the bench workload itself (L=7): DRAM traffic per launch for roofline.traffic
timeout 600 ncu --set full --clock-control none -k regex:k_laplace_apply --launch-skip 3 -c 1 -f -o $O/${TAG}_vmult_L7 \
  python tools/time_vmult.py --nref 7 --degree 4 --reps 2 > $O/${TAG}_ncu_full_L7.log 2>&1; echo "ncu full L7 rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_laplace_q1_slab --launch-skip 3 -c 1 -f -o $O/${TAG}_q1_slab_L6 \
  python tools/time_vmult.py --nref 6 --degree 4 --deform smooth --reps 2 > $O/${TAG}_ncu_q1.log 2>&1; echo "ncu q1 rc=$?"
for L in 6 7; do timeout 120 python tools/time_transfer.py --nref $L >> $O/${TAG}_time_transfer.txt 2>&1; done
ls -la $O | tail -20
