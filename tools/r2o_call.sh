#!/bin/bash
# r2o: the fused peer exchange as the default on 4 GPUs: parity over all hierarchies and partition policies, bench both ways
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
MGB_VERBOSE=1 timeout 600 $TR --master-port 29631 tools/check_multigpu.py > $O/r2o_check_4gpu.log 2>&1; echo "parity rc=$? $(grep PARITY $O/r2o_check_4gpu.log) $(grep -c 'peer-memory ghost exchange enabled' $O/r2o_check_4gpu.log) problems with peer exchange"
grep -E "FAIL|rror" $O/r2o_check_4gpu.log | head -5
i=0
for env in "MGB_P2P=0" "MGB_P2P=1"; do
  i=$((i+1))
  env $env timeout 400 $TR --master-port 2964$i bench.py --gpus 4 --steps 10 --warmup 3 2>/dev/null | grep '^{' > $O/r2o_bench_4gpu_$i.json
  echo "$env: $(python -c "import json;d=json.load(open('$O/r2o_bench_4gpu_$i.json'));print(d['ms_per_step'], d['value']/1e9, d['gpu_launches'])")"
done
