#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
VO2W_PAIR_PROF=1 VO2W_PAIR_TRACE=1 timeout 300 python tools/profile_target.py 1 64 > gpurun_out/trace.log 2>&1; echo "rc=$?"
