#!/bin/bash
# r2k: config 5 at size -- hypercube L=8 p=4 (1 076 890 625 DoFs) on 8 B200, weak-scaling point of the L=7 single-GPU bench,
# then the strong-scaling line of the default workload on the same box for reference
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
nproc; free -g | head -2
( while true; do free -g | awk 'NR==2{print "mem used GB", $3}'; sleep 20; done ) > $O/r2k_mem.log 2>&1 &
MEMPID=$!
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 8 --steps 5 --warmup 3 --scaling weak \
  > $O/r2k_bench_weak_8gpu.json 2> $O/r2k_bench_weak_8gpu.err; echo "weak rc=$?"; cut -c1-700 $O/r2k_bench_weak_8gpu.json; tail -5 $O/r2k_bench_weak_8gpu.err | cut -c1-300
kill $MEMPID; sort -k4 -n $O/r2k_mem.log | tail -1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29562 bench.py --gpus 8 --steps 5 --warmup 3 \
  > $O/r2k_bench_strong_8gpu.json 2> $O/r2k_bench_strong_8gpu.err; echo "strong rc=$?"; cut -c1-400 $O/r2k_bench_strong_8gpu.json
