#!/bin/bash
# r2final3: reference arm (host cores of the GPU box) with the final C restatement: 8 cells per vector, even-odd sweeps, block colouring
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 600 python bench.py --impl reference > $O/r2final3_bench_ref.json 2> $O/r2final3_bench_ref.err; echo "ref rc=$?"
python -c "import json;d=json.load(open('$O/r2final3_bench_ref.json'));print(d['value']/1e6, d['ms_per_step'], d['cpu_baseline']['cores'], d['n_iterations'])"
