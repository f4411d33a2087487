#!/bin/bash
# one process per section so a trapped kernel cannot poison the others
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
for s in lr simt umma gen_fp32 gen_bf16; do
  timeout 300 python tools/gpu_debug.py $s > gpurun_out/debug_$s.log 2>&1
  echo "section $s exit $?" >> gpurun_out/debug_summary.log
done
cat gpurun_out/debug_summary.log
tail -n 40 gpurun_out/debug_umma.log gpurun_out/debug_gen_bf16.log
