#!/bin/bash
# usage (under gpurun): bash tools/run_probe.sh [kernels...]
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv | tee gpurun_out/probe.log
for k in "${@:-splice ln gemm patch attn tower}"; do
  for kk in $k; do
    echo "=== $kk" | tee -a gpurun_out/probe.log
    timeout -k 5 150 python tools/gpu_probe.py $kk 2>&1 | tail -60 | tee -a gpurun_out/probe.log
    echo "exit ${PIPESTATUS[0]}" | tee -a gpurun_out/probe.log
  done
done
