#!/bin/bash
# r2h: GPU suite with the new tests (chunked loop, composed hierarchies, Gaussian problem, float oracle), bench lines
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q > $O/r2h_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 $O/r2h_pytest.log)"
grep -E "FAILED|ERROR" $O/r2h_pytest.log | head -20
timeout 900 python bench.py --steps 5 --warmup 3 > $O/r2h_bench_L7.json 2> $O/r2h_bench_L7.err; echo "bench rc=$?"; cut -c1-400 $O/r2h_bench_L7.json; tail -3 $O/r2h_bench_L7.err
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > $O/r2h_bench_ref.json 2> $O/r2h_bench_ref.err; echo "ref rc=$?"; cut -c1-900 $O/r2h_bench_ref.json; tail -3 $O/r2h_bench_ref.err
nproc; free -g | head -2
