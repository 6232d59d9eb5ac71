#!/bin/bash
# NVLink evidence for the peer put/get kernels (2 GPUs visible, ONE process, one GPU launching).
# 1. plain run (must exit 0), 2. NVLink/DRAM byte counters per kernel, 3. one full-set capture of the peer put.
set -u
mkdir -p gpurun_out
build/bin/nvl_prof > gpurun_out/nvl_prof_plain.log 2>&1 || { echo "nvl_prof failed"; tail -5 gpurun_out/nvl_prof_plain.log; exit 1; }
M=nvltx__bytes.sum,nvlrx__bytes.sum,nvltx__bytes_data_user.sum,nvlrx__bytes_data_user.sum,nvltx__bytes_data_protocol.sum,nvlrx__bytes_data_protocol.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum
timeout 600 ncu --clock-control none --metrics $M --csv --log-file gpurun_out/nvl_prof_metrics.csv build/bin/nvl_prof > gpurun_out/nvl_prof_ncu.log 2>&1
echo "ncu metrics rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ocm_copy_v32 -s 2 -c 1 -o gpurun_out/prof_put_peer build/bin/nvl_prof > gpurun_out/nvl_prof_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out | tail -8
