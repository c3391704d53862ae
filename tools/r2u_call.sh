#!/bin/bash
# r2u: dst = 0 of the next Chebyshev product cleared inside the cell loop (k_laplace_apply<P,T,true>): GPU suite, then the
# bench solve with the in-loop zeroing (default) and with a memset before every cell loop (MGB_ZERO_IN_CELL_LOOP=0)
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/r2u_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 $O/r2u_pytest.log)"
grep -E "^FAILED|^E  " $O/r2u_pytest.log | head -8
for v in 1 0 1 0; do
  MGB_ZERO_IN_CELL_LOOP=$v timeout 600 python bench.py --no-cpu > $O/r2u_bench_zero$v.json 2> $O/r2u_bench_zero$v.err; echo "bench zero=$v rc=$?"
  python - <<PY
import json
d=json.load(open("$O/r2u_bench_zero$v.json"))
print("zero_in_loop=$v", "ms_per_step", round(d["ms_per_step"],3), "GDoF/s", round(d["value"]/1e9,3), "vmult avg ms", round(d["roofline"]["avg_launch_ms"],4), "frac", round(d["roofline"]["frac"],4), "launches", d["gpu_launches"], d["clocks"]["sm_mhz"])
PY
done
