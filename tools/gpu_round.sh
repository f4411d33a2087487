#!/bin/bash
# tests + smoke + bench, then the ncu passes (launch list, one full capture) on a short target
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 5 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/smoke.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -n 1 gpurun_out/bench.log | cut -c1-2500
if [ "$1" == "ncu" ]; then
B=${2:-64}
timeout 300 python tools/profile_target.py 2 $B > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python tools/profile_target.py 2 $B > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 300 python tools/profile_target.py 2 $B > gpurun_out/plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:pair_umma -s 28 -c 28 -o gpurun_out/prof -f python tools/profile_target.py 2 $B > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
# the report itself can exceed what gpurun copies back: keep the raw page as CSV
ncu -i gpurun_out/prof.ncu-rep --page raw --csv > gpurun_out/prof_raw.csv 2>/dev/null; rm -f gpurun_out/prof.ncu-rep
fi
