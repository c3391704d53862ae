#!/bin/bash
# r2x: final tree on 2 GPUs: multi-rank parity tests (all exchange variants), tools/check_multigpu.py, bench lines of both arms
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_gpu_multirank.py -m gpu -q > $O/r2x_pytest_2gpu.log 2>&1; echo "pytest rc=$? $(tail -1 $O/r2x_pytest_2gpu.log)"
grep -E "FAILED|Error|trap" $O/r2x_pytest_2gpu.log | head -5
timeout 400 $TR --master-port 29571 bench.py --gpus 2 --steps 5 --warmup 3 2> $O/r2x_bench_2gpu.err | grep '^{' > $O/r2x_bench_2gpu.json
python -c "import json;d=json.load(open('$O/r2x_bench_2gpu.json'));print('2 GPUs', d['ms_per_step'], d['value']/1e9, d['e2e']['value']/1e9, d['gpu_launches'], d.get('parity_check'))"
timeout 600 $TR --master-port 29572 bench.py --impl reference --gpus 2 --steps 5 --warmup 3 2> $O/r2x_ref_2gpu.err | grep '^{' > $O/r2x_ref_2gpu.json
python -c "import json;d=json.load(open('$O/r2x_ref_2gpu.json'));print('ref arm under torchrun', d['value']/1e6, d['ms_per_step'], d['cpu_baseline']['cores'])"
CUDA_VISIBLE_DEVICES=0 timeout 400 python bench.py --number float --no-cpu > $O/r2x_bench_L7_float.json 2> $O/r2x_bench_L7_float.err
python -c "import json;d=json.load(open('$O/r2x_bench_L7_float.json'));print('float levels', d['ms_per_step'], d['value']/1e9, d['roofline']['kernel'], d['roofline']['frac'])"
