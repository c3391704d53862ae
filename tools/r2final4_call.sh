#!/bin/bash
# r2final4: the checkers that use the C restatement, after its block colouring / even-odd sweeps: scale tests and the full bench line
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_scale.py -m gpu -x -q > $O/r2final4_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 $O/r2final4_pytest.log)"
timeout 400 python bench.py > $O/r2final4_bench_L7.json 2> $O/r2final4_bench_L7.err; echo "bench rc=$?"
python -c "import json;d=json.load(open('$O/r2final4_bench_L7.json'));print(d['ms_per_step'], d['value']/1e9, d['cpu_baseline']['value']/1e6, d['parity_check'])"
