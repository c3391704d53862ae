#!/bin/bash
# r2m: fused single-kernel peer exchange on 2 GPUs: parity, per-exchange time, bench lines
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_gpu_multirank.py -m gpu -q > $O/r2m_pytest_2gpu.log 2>&1; echo "pytest rc=$? $(tail -1 $O/r2m_pytest_2gpu.log)"
grep -E "FAILED|Error|trap" $O/r2m_pytest_2gpu.log | head -5
for f in 1 0; do for L in 4 6; do
  echo "MGB_P2P_FUSED=$f" >> $O/r2m_time_halo.txt
  MGB_P2P_FUSED=$f timeout 200 $TR --master-port 2957$L tools/time_halo.py --nref $L 2>/dev/null | grep '^{' >> $O/r2m_time_halo.txt
done; done
cat $O/r2m_time_halo.txt
i=0
for env in "MGB_P2P=0" "MGB_P2P=1" "MGB_P2P=1 MGB_P2P_FUSED=0"; do
  i=$((i+1))
  env $env timeout 400 $TR --master-port 2958$i bench.py --gpus 2 --steps 5 --warmup 3 2>/dev/null | grep '^{' > $O/r2m_bench_2gpu_$i.json
  echo "$env: $(python -c "import json;d=json.load(open('$O/r2m_bench_2gpu_$i.json'));print(d['ms_per_step'], d['value']/1e9, d['gpu_launches'])")"
done
