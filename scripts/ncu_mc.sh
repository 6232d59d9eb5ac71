#!/bin/bash
# NVLink counters of the NVLS kernels (multimem.st put, multimem.ld_reduce get): one process, 2 GPUs.
set -u
mkdir -p gpurun_out
build/bin/mc_probe > gpurun_out/mc_probe_plain.jsonl 2>&1 || { echo "mc_probe failed"; exit 1; }
cat gpurun_out/mc_probe_plain.jsonl
grep -q '"ok": true' gpurun_out/mc_probe_plain.jsonl || exit 0
M=nvltx__bytes.sum,nvlrx__bytes.sum,nvltx__bytes_data_user.sum,nvlrx__bytes_data_user.sum,nvltx__bytes_data_protocol.sum,nvlrx__bytes_data_protocol.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum
timeout 300 ncu --clock-control none --metrics $M -k regex:multimem\|reduce --csv --log-file gpurun_out/mc_probe_metrics.csv build/bin/mc_probe > gpurun_out/mc_probe_ncu.log 2>&1
echo "ncu rc=$?"
