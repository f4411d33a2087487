#!/bin/bash
# Round-2 profile evidence: (a) launch list of bench.py itself, (b) --set full over the 28 fused launches of one forward
# at the bench sub-batch, (c) --set full over the five stream_umma_kernel launches.  Each ncu pass follows a plain run
# of the same command that exited 0.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
BENCH="python bench.py --steps 2 --warmup 1 --no-extras --no-cpu-baseline"
timeout 600 $BENCH > gpurun_out/bench_plain.log 2> gpurun_out/bench_plain.err && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_bench_launches.csv $BENCH > gpurun_out/ncu_bench.log 2>&1
echo "ncu bench launches rc=$?"
timeout 300 python tools/profile_target.py 2 128 > gpurun_out/plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:pair_umma -s 28 -c 28 -o gpurun_out/pair_prof -f python tools/profile_target.py 2 128 > gpurun_out/ncu_pair.log 2>&1
echo "ncu pair full rc=$?"
ncu -i gpurun_out/pair_prof.ncu-rep --page raw --csv > gpurun_out/r02_pair_umma_full.csv 2>/dev/null; rm -f gpurun_out/pair_prof.ncu-rep
timeout 300 python tools/profile_target.py 2 128 > gpurun_out/plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stream_umma -s 5 -c 5 -o gpurun_out/su_prof -f python tools/profile_target.py 2 128 > gpurun_out/ncu_su.log 2>&1
echo "ncu stream full rc=$?"
ncu -i gpurun_out/su_prof.ncu-rep --page raw --csv > gpurun_out/r02_stream_umma_full.csv 2>/dev/null; rm -f gpurun_out/su_prof.ncu-rep
ls -la gpurun_out/r02_*
