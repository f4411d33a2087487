#!/bin/bash
# tests + bench + per-role cycle accounting of the fused kernel (no ncu)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 8 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/smoke.log
timeout 600 python bench.py --steps 3 --warmup 3 ${BENCH_ARGS} > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -n 1 gpurun_out/bench.log | cut -c1-3000
tail -n 5 gpurun_out/bench.err
VO2W_PAIR_PROF=1 timeout 300 python tools/profile_target.py 1 64 > gpurun_out/pairprof.log 2>&1; echo "prof rc=$?"
grep "pair prof" gpurun_out/pairprof.log | cut -c1-400
