#!/bin/bash
# One 8-GPU lease: headline bench at N=8, BASELINE configs 3/4/5 at scale.  Everything under `timeout`; the most
# important results first, the biggest sort last.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 240 $TR --nproc-per-node 8 --master-port 29701 bench.py --gpus 8 --steps 20 --warmup 5 > $O/bench8_r2.json 2> $O/bench8_r2.err
echo "bench8 rc=$?"
# config 4, mid scale: 2^30 keys per rank = 64 GiB in all, a quarter of runs + output on the pinned host tier
timeout 200 $TR --nproc-per-node 8 --master-port 29704 bench/workloads.py --only oocsort --sort-keys $((1<<30)) --sort-chunk $((1<<27)) \
    --hbm-frac 0.75 --gpu-bytes 40G --host-bytes 8G > $O/oocsort8_64g.jsonl 2> $O/oocsort8_64g.err
echo "oocsort 64g rc=$?"
# config 5: 64 clients x 10k alloc/free of <= 4 KiB remote-GPU pairs while every rank runs the strided fp8 put loop
timeout 240 python bench/stress.py --apps 64 --iters 10000 --ranks 8 --kind 7 --max-bytes 4096 --gpu --puts > $O/stress8_r2.json 2> $O/stress8_r2.err
echo "stress rc=$?"
# config 3 (GUPS over the pooled table) + sort with both finals
timeout 240 $TR --nproc-per-node 8 --master-port 29703 bench/workloads.py --sort-keys $((1<<27)) --sort-chunk $((1<<24)) --sort-final both > $O/workloads8_r2.jsonl 2> $O/workloads8_r2.err
echo "workloads rc=$?"
# config 4 at scale: 2^32 keys per rank = 256 GiB in all
timeout 330 $TR --nproc-per-node 8 --master-port 29705 bench/workloads.py --only oocsort --sort-keys $((1<<32)) --sort-chunk $((1<<28)) \
    --hbm-frac 0.75 --gpu-bytes 100G --host-bytes 24G > $O/oocsort8_256g.jsonl 2> $O/oocsort8_256g.err
echo "oocsort 256g rc=$?"
tail -c 300 $O/bench8_r2.err $O/workloads8_r2.err $O/oocsort8_64g.err $O/oocsort8_256g.err $O/stress8_r2.err
cat $O/oocsort8_64g.jsonl $O/oocsort8_256g.jsonl $O/stress8_r2.json | cut -c1-500
