#!/bin/bash
# r2final8: the final round-2 tree on 8 B200: strong-scaling line of the bench workload, then the weak-scaling point (L=8, 1.08e9 DoFs)
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29562 bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu > $O/r2final_bench_strong_8gpu.json 2> $O/r2final_bench_strong_8gpu.err; echo "strong rc=$?"
python -c "import json;d=json.load(open('$O/r2final_bench_strong_8gpu.json'));print('strong', d['ms_per_step'], d['value']/1e9, d['e2e']['value']/1e9)"
timeout 260 $TR --master-port 29561 bench.py --gpus 8 --steps 5 --warmup 3 --scaling weak --no-cpu > $O/r2final_bench_weak_8gpu.json 2> $O/r2final_bench_weak_8gpu.err; echo "weak rc=$?"
python -c "import json;d=json.load(open('$O/r2final_bench_weak_8gpu.json'));print('weak', d['ms_per_step'], d['value']/1e9, d['config']['n_dofs'])"
