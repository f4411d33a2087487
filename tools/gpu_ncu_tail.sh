#!/bin/bash
# ncu evidence for the acoustic tail: launch list of one pass of tools/bench_tail.py and a
# full capture of the tensor-pipe convolution (conv_tf32.cu) launches of one decoder layer + PostNet
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python tools/bench_tail.py --steps 1 --warmup 1 > gpurun_out/tail_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/tail_launches.csv python tools/bench_tail.py --steps 1 --warmup 1 > gpurun_out/tail_ncu.log 2>&1
echo "ncu tail launches rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:tf32_conv -s 24 -c 9 -o gpurun_out/tail_prof -f python tools/bench_tail.py --steps 1 --warmup 1 > gpurun_out/tail_ncu_full.log 2>&1
echo "ncu tail full rc=$?"
ncu -i gpurun_out/tail_prof.ncu-rep --page raw --csv > gpurun_out/tail_prof_raw.csv 2>/dev/null; rm -f gpurun_out/tail_prof.ncu-rep
