#!/bin/bash
# Second 8-GPU lease of round 2: out-of-core sort at 512 GiB (fallback 256 GiB) with the VMM host tier, and the
# allocation stress with node-local daemon fulfilment.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
free -g | head -2
timeout 170 $TR --nproc-per-node 8 --master-port 29705 bench/workloads.py --only oocsort --sort-keys $((1<<33)) --sort-chunk $((1<<28)) \
    --sort-repeat 2 --hbm-frac 0.75 --gpu-bytes 150G --host-bytes 40G > $O/oocsort8_512g.jsonl 2> $O/oocsort8_512g.err
rc=$?; echo "oocsort 512g rc=$rc"
if [ $rc -ne 0 ]; then
  grep "ocm E\|Error\|error" $O/oocsort8_512g.err | head -5
  timeout 120 $TR --nproc-per-node 8 --master-port 29706 bench/workloads.py --only oocsort --sort-keys $((1<<32)) --sort-chunk $((1<<28)) \
      --sort-repeat 2 --hbm-frac 0.75 --gpu-bytes 100G --host-bytes 24G > $O/oocsort8_256g_vmm.jsonl 2> $O/oocsort8_256g_vmm.err
  echo "oocsort 256g rc=$?"
fi
timeout 110 python bench/stress.py --apps 64 --iters 3000 --ranks 8 --kind 7 --max-bytes 4096 --gpu --puts > $O/stress8_r2b.json 2> $O/stress8_r2b.err
echo "stress rc=$?"
cat $O/oocsort8_512g.jsonl $O/oocsort8_256g_vmm.jsonl 2>/dev/null | cut -c1-900
cut -c1-700 $O/stress8_r2b.json; tail -c 300 $O/stress8_r2b.err
