#!/bin/bash
# r2p: batch host-buffer solves (pipelined copies), full GPU suite, bench line
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_host_batch.py -m gpu -q > $O/r2p_pytest_batch.log 2>&1; echo "batch rc=$? $(tail -1 $O/r2p_pytest_batch.log)"
grep -E "^E " $O/r2p_pytest_batch.log | head -5
timeout 900 python bench.py --steps 5 --warmup 3 > $O/r2p_bench_L7.json 2> $O/r2p_bench_L7.err; echo "bench rc=$?"; tail -2 $O/r2p_bench_L7.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2p_bench_L7.json'))
print("value", d["value"]/1e9, "ms", d["ms_per_step"], "e2e", d["e2e"]["value"]/1e9, d["e2e"]["ms_per_step"], "single", d["e2e"]["one_call_per_step"]["value"]/1e9, d["e2e"]["one_call_per_step"]["ms_per_step"])
print("roofline", d["roofline"]["frac"], d["roofline"]["fp64"]["achieved_tflops"], d["roofline"]["fp64"]["frac"], "parity", d["parity_check"]["ok"], "cpu", d["cpu_baseline"]["value"]/1e6)
PY
