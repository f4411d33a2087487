#!/bin/bash
# Round-2 profile evidence outside the bf16 hot path: (a) launch list of one pass of every stage of the acoustic tail
# (tools/bench_tail.py at B=64), (b) launch list of one fp32-mode Generator forward at B=64 x T=172, (c) --set full over
# the first 40 tf32_conv_kernel launches of that forward.  Each ncu pass follows a plain run of the same command that
# exited 0.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python tools/bench_tail.py --steps 1 --warmup 1 > gpurun_out/tail_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_tail_launches_b64.csv python tools/bench_tail.py --steps 1 --warmup 1 > gpurun_out/tail_ncu.log 2>&1
echo "ncu tail launches rc=$?"
timeout 300 python tools/profile_target.py 1 64 fp32 > gpurun_out/fp32_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_fp32_launches_b64.csv python tools/profile_target.py 1 64 fp32 > gpurun_out/fp32_ncu.log 2>&1
echo "ncu fp32 launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tf32_conv -c 40 -o gpurun_out/fp32_prof -f python tools/profile_target.py 1 64 fp32 > gpurun_out/fp32_ncu_full.log 2>&1
echo "ncu fp32 full rc=$?"
ncu -i gpurun_out/fp32_prof.ncu-rep --page raw --csv > gpurun_out/r02_fp32_tf32_conv_full.csv 2>/dev/null; rm -f gpurun_out/fp32_prof.ncu-rep
tail -2 gpurun_out/tail_plain.log gpurun_out/fp32_plain.log
ls -la gpurun_out/r02_tail_* gpurun_out/r02_fp32_*
