#!/bin/bash
# r2n: fused peer exchange on 8 GPUs: per-exchange time against NCCL, parity, strong-scaling bench both ways
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
for L in 5 6 7; do
  timeout 200 $TR --master-port 2960$L tools/time_halo.py --nref $L 2>/dev/null | grep '^{' >> $O/r2n_time_halo_8gpu.txt
done
cat $O/r2n_time_halo_8gpu.txt
MGB_P2P=1 MGB_CHECK_ONLY=DefaultPolicy timeout 400 $TR --master-port 29611 tools/check_multigpu.py > $O/r2n_check_8gpu_p2p.log 2>&1; echo "parity rc=$? $(grep PARITY $O/r2n_check_8gpu_p2p.log)"
i=0
for env in "MGB_P2P=0" "MGB_P2P=1"; do
  i=$((i+1))
  env $env timeout 400 $TR --master-port 2962$i bench.py --gpus 8 --steps 10 --warmup 3 2>/dev/null | grep '^{' > $O/r2n_bench_8gpu_$i.json
  echo "$env: $(python -c "import json;d=json.load(open('$O/r2n_bench_8gpu_$i.json'));print(d['ms_per_step'], d['value']/1e9, d['gpu_launches'], d['roofline']['avg_launch_ms'])")"
done
