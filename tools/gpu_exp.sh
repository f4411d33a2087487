#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
TAG=new timeout 200 python tools/launch_times.py 64 2>&1 | grep -v Removing | tee gpurun_out/exp.log
TAG=b16 timeout 200 python tools/launch_times.py 16 2>&1 | grep -v Removing | tee -a gpurun_out/exp.log
