#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for d in 0; do TAG=dbg$d VO2W_PAIR_DBG=$d timeout 200 python tools/launch_times.py 64 2>&1 | grep -v Removing; done | tee gpurun_out/exp.log
