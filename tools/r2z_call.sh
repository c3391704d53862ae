#!/bin/bash
# r2z: secondary (HBM-bound) kernels at the bench size: event timings, then ncu --set full of the Chebyshev updates and the transfers
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 600 python tools/time_secondary.py > $O/r2z_time_secondary.txt 2> $O/r2z_time_secondary.err; echo "time rc=$?"; cut -c1-260 $O/r2z_time_secondary.txt
timeout 600 ncu --set full --clock-control none -k regex:k_cheb -c 3 -f -o $O/r2z_cheb python tools/time_secondary.py --reps 1 > $O/r2z_ncu_cheb.log 2>&1; echo "ncu cheb rc=$?"
timeout 600 ncu --set full --clock-control none -k regex:k_transfer -c 2 -f -o $O/r2z_transfer python tools/time_secondary.py --reps 1 > $O/r2z_ncu_transfer.log 2>&1; echo "ncu transfer rc=$?"
ls -la $O/r2z_*.ncu-rep
