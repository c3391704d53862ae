#!/bin/bash
# r2b: first GPU pass of the chunked cell loop (device/chunk.cuh): parity tests, vmult timings against the atomic kernel
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/r2b_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 $O/r2b_pytest.log)"
for v in 1 0; do
  for cfg in "hypercube 7 4" "hypercube 7 3" "hypercube 7 2" "hypercube 7 1" "quadrant 7 4" "hypercube 6 5"; do
    set -- $cfg
    echo "MGB_CHUNKED=$v $cfg" >> $O/r2b_time_vmult.txt
    MGB_CHUNKED=$v timeout 200 python tools/time_vmult.py --geometry $1 --nref $2 --degree $3 --reps 30 >> $O/r2b_time_vmult.txt 2>&1
  done
done
MGB_CHUNKED=1 timeout 200 python tools/time_vmult.py --nref 7 --degree 4 --number float --reps 30 >> $O/r2b_time_vmult.txt 2>&1
cat $O/r2b_time_vmult.txt
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $O/r2b_bench_L7.json 2> $O/r2b_bench_L7.err; echo "bench rc=$?"; cat $O/r2b_bench_L7.json; tail -3 $O/r2b_bench_L7.err
MGB_FUSE_CHEB=0 timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $O/r2b_bench_L7_nofuse.json 2>/dev/null; cat $O/r2b_bench_L7_nofuse.json
MGB_CHUNKED=0 timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $O/r2b_bench_L7_legacy.json 2>/dev/null; cat $O/r2b_bench_L7_legacy.json
