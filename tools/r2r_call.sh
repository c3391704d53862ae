#!/bin/bash
# r2r: round-2 evidence: full GPU suite, bench line, ncu launch list of the same command, ncu --set full of the dominant kernel
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/r2r_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 $O/r2r_pytest.log)"
grep -E "^FAILED|^E  " $O/r2r_pytest.log | head -8
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2r_smoke.log 2>&1; echo "smoke rc=$? $(tail -1 $O/r2r_smoke.log)"
timeout 900 python bench.py > $O/r2r_bench_L7.json 2> $O/r2r_bench_L7.err; echo "bench rc=$?"; tail -2 $O/r2r_bench_L7.err
timeout 300 python tools/time_vmult.py --geometry quadrant --nref 7 --degree 4 --reps 50 > $O/r2r_time_vmult_quadrant.txt 2>&1; tail -1 $O/r2r_time_vmult_quadrant.txt | cut -c1-330
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu > $O/r2r_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/r2r_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > $O/r2r_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 300 python tools/time_vmult.py --nref 7 --degree 4 --reps 3 > $O/r2r_plain2.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_laplace_apply --launch-skip 3 -c 1 -f -o $O/r2r_vmult_L7 python tools/time_vmult.py --nref 7 --degree 4 --reps 3 > $O/r2r_ncu_full.log 2>&1; echo "ncu full rc=$?"
