#!/bin/bash
# r2c: second pass of the chunked cell loop (compact tables, unrolled epilogue loops, 3 blocks per SM)
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/r2c_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 $O/r2c_pytest.log)"
for cfg in "hypercube 7 4" "hypercube 7 3" "hypercube 7 2" "hypercube 7 1" "quadrant 7 4"; do
  set -- $cfg
  echo "chunked $cfg" >> $O/r2c_time_vmult.txt
  timeout 200 python tools/time_vmult.py --geometry $1 --nref $2 --degree $3 --reps 30 >> $O/r2c_time_vmult.txt 2>&1
done
echo "chunked minb2 hypercube 7 4" >> $O/r2c_time_vmult.txt
MGB200_LIB=$PWD/variants/libmgb200_minb2.so timeout 200 python tools/time_vmult.py --nref 7 --degree 4 --reps 30 >> $O/r2c_time_vmult.txt 2>&1
cat $O/r2c_time_vmult.txt
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $O/r2c_bench_L7.json 2> $O/r2c_bench_L7.err; echo "bench rc=$?"; cat $O/r2c_bench_L7.json
timeout 200 python tools/time_vmult.py --nref 6 --degree 4 --reps 3 > $O/r2c_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_laplace_chunks --launch-skip 3 -c 1 -f -o $O/r2c_chunks_L6 \
  python tools/time_vmult.py --nref 6 --degree 4 --reps 3 > $O/r2c_ncu.log 2>&1; echo "ncu rc=$?"
