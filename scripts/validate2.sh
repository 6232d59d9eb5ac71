TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for frac in 0.5 1.0; do
timeout 200 $TR --nproc-per-node 2 --master-port 29711 bench/workloads.py --only oocsort --sort-keys $((1<<27)) --sort-chunk $((1<<24)) --sort-repeat 2 --hbm-frac $frac --host-bytes 4G > gpurun_out/oocsort2_$frac.jsonl 2> gpurun_out/oocsort2_$frac.err; echo "rc=$?"
cut -c1-900 gpurun_out/oocsort2_$frac.jsonl
done
