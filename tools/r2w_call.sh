#!/bin/bash
# r2w: full-size property tests (L=7, 135 M DoFs) + whole GPU suite + smoke
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_full_size.py -m gpu -x -q --durations=5 > $O/r2w_full_size.log 2>&1; echo "full size rc=$? $(tail -1 $O/r2w_full_size.log)"
grep -E "^FAILED|^E  |s call|s setup" $O/r2w_full_size.log | head -12
timeout 1200 python -m pytest tests -m gpu -x -q > $O/r2w_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 $O/r2w_pytest.log)"
