#!/bin/bash
# tests + smoke + bench, then the ncu passes (launch list, one full capture) on a short target
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 5 gpurun_out/pytest_gpu.log
python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/smoke.log
python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -n 1 gpurun_out/bench.log | cut -c1-1500
if [ "$1" == "ncu" ]; then
python tools/profile_target.py 2 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 72 -c 71 --csv --log-file gpurun_out/launches.csv python tools/profile_target.py 2 > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
python tools/profile_target.py 2 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:umma_conv -s 89 -c 6 -o gpurun_out/prof -f python tools/profile_target.py 2 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
fi
