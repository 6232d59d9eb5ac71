set -o pipefail
python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/pytest_gpu2b.log; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu2b.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 200 $TR --nproc-per-node 2 --master-port 29711 bench/workloads.py --sort-keys $((1<<27)) --sort-chunk $((1<<24)) --sort-repeat 2 --host-bytes 4G > gpurun_out/workloads2_r2.jsonl 2> gpurun_out/workloads2_r2.err; echo "workloads rc=$?"
timeout 200 python bench/stress.py --apps 16 --iters 3000 --ranks 2 --kind 7 --max-bytes 4096 --gpu --puts > gpurun_out/stress2_r2.json 2> gpurun_out/stress2_r2.err; echo "stress rc=$?"
cut -c1-400 gpurun_out/workloads2_r2.jsonl; cut -c1-900 gpurun_out/stress2_r2.json; tail -c 300 gpurun_out/stress2_r2.err
