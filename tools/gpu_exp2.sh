#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -q -x -k "single_layer or stream_kernel or generator" > gpurun_out/t_k.log 2>&1; echo "kernel tests rc=$?"; tail -n 3 gpurun_out/t_k.log
run() { TAG="$1" timeout 300 python tools/launch_times.py $2 2>&1 | grep -E "^tag|^us:" | sed -E 's/ 2:[0-9]+//g'; }
run b128 128
run b256 256
run b128 128
run b256 256
run b64 64
