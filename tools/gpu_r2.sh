#!/bin/bash
# Round-2 GPU pass: new-kernel known-answer tests first (own processes), then the whole GPU suite, smoke, the parity
# statistics, per-launch times (new single-conv path vs the old one) and the bench.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_parity.py -q -k "single_layer or stream_kernel" > gpurun_out/t_stream.log 2>&1; echo "stream tests rc=$?"; tail -n 4 gpurun_out/t_stream.log
timeout 600 python -m pytest tests/test_gpu_parity.py -q -k "conv_post_on_board or fused_resblock" > gpurun_out/t_post.log 2>&1; echo "post tests rc=$?"; tail -n 4 gpurun_out/t_post.log
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 15 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/smoke.log
timeout 600 python tools/gpu_stats.py 16 > gpurun_out/stats.log 2>&1; echo "stats rc=$?"; tail -n 30 gpurun_out/stats.log
TAG=new timeout 300 python tools/launch_times.py 128 > gpurun_out/lt_new.log 2>&1; tail -n 2 gpurun_out/lt_new.log
TAG=old VO2W_SINGLE_CONV=old VO2W_FUSE_POST=0 timeout 300 python tools/launch_times.py 128 > gpurun_out/lt_old.log 2>&1; tail -n 2 gpurun_out/lt_old.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -n 1 gpurun_out/bench.log | cut -c1-6000
tail -n 5 gpurun_out/bench.err
