#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for l in A B A B; do TAG=lib$l VO2W_LIB=$PWD/tools/lib$l.so timeout 200 python tools/launch_times.py 64 2>&1 | grep -v Removing; done | tee gpurun_out/exp.log
