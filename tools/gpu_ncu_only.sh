#!/bin/bash
# the two ncu passes of gpu_round.sh without tests / bench
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
B=${1:-64}
timeout 300 python tools/profile_target.py 2 $B > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python tools/profile_target.py 2 $B > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 300 python tools/profile_target.py 2 $B > gpurun_out/plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:pair_umma -s 28 -c 28 -o gpurun_out/prof -f python tools/profile_target.py 2 $B > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
ncu -i gpurun_out/prof.ncu-rep --page raw --csv > gpurun_out/prof_raw.csv 2>/dev/null; rm -f gpurun_out/prof.ncu-rep
