set -o pipefail
python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > gpurun_out/pytest_gpu2d.log; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu2d.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 200 $TR --nproc-per-node 2 --master-port 29711 bench/workloads.py --only oocsort --sort-keys $((1<<27)) --sort-chunk $((1<<24)) --sort-repeat 2 --hbm-frac 0.5 --host-bytes 4G > gpurun_out/oocsort2_vmm.jsonl 2> gpurun_out/oocsort2_vmm.err; echo "oocsort rc=$?"
cut -c1-900 gpurun_out/oocsort2_vmm.jsonl; grep "ocm E\|ocm W" gpurun_out/oocsort2_vmm.err | head -5
timeout 200 $TR --nproc-per-node 2 --master-port 29712 bench.py --gpus 2 --steps 10 --warmup 3 --curve-max $((1<<24)) > gpurun_out/bench2_vmm.json 2> gpurun_out/bench2_vmm.err; echo "bench rc=$?"
cut -c1-1300 gpurun_out/bench2_vmm.json
true
