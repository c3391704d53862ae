#!/bin/bash
# r2final: last validation of the round-2 tree:
# full GPU suite, smoke, bench line (GPU arm) and the reference arm (host cores, batched C restatement)
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/r2final_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 $O/r2final_pytest.log)"
grep -E "^FAILED|^E  " $O/r2final_pytest.log | head -8
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2final_smoke.log 2>&1; echo "smoke rc=$? $(tail -1 $O/r2final_smoke.log)"
timeout 900 python bench.py > $O/r2final_bench_L7.json 2> $O/r2final_bench_L7.err; echo "bench rc=$?"; tail -2 $O/r2final_bench_L7.err
timeout 900 python bench.py --impl reference > $O/r2final_bench_ref.json 2> $O/r2final_bench_ref.err; echo "ref rc=$?"; cut -c1-400 $O/r2final_bench_ref.json
