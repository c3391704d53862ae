#!/bin/bash
# r2final2: the tree after the host-setup changes (threads, huge pages): GPU suite and the bench line (setup_s in config)
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/r2final2_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 $O/r2final2_pytest.log)"
grep -E "^FAILED|^E  " $O/r2final2_pytest.log | head -8
MGB_SETUP_TIMING=1 timeout 400 python bench.py --no-cpu > $O/r2final2_bench_L7.json 2> $O/r2final2_bench_L7.err; echo "bench rc=$?"
grep "DoFMap::build 2097152" $O/r2final2_bench_L7.err | head -2
python -c "import json;d=json.load(open('$O/r2final2_bench_L7.json'));print(d['ms_per_step'], d['value']/1e9, d['config']['setup_s'])"
