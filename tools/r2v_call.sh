#!/bin/bash
# r2v: kinds of store / place in the batch for the side zeroing of k_laplace_apply<P,T,true> (variants/, tools/build_variants.sh)
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
run() { # tag, lib, env
  MGB200_LIB=$2 MGB_ZERO_IN_CELL_LOOP=$3 timeout 600 python bench.py --no-cpu --steps 4 > $O/r2v_$1.json 2> $O/r2v_$1.err || { echo "$1 failed"; tail -3 $O/r2v_$1.err; return; }
  python - <<PY
import json
d=json.load(open("$O/r2v_$1.json"))
print("$1", "ms_per_step", round(d["ms_per_step"],3), "vmult avg ms", round(d["roofline"]["avg_launch_ms"],4), "parity", d.get("parity_check",{}).get("ok"), d["clocks"]["sm_mhz"])
PY
}
run memset "" 0
run plain "" 1
for t in zcs zwt zcs_end z_end; do run $t $PWD/variants/libmgb200_$t.so 1; done
