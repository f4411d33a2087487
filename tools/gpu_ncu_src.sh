#!/bin/bash
# ncu --set full with source correlation for ONE fused launch (index $1 among the pair_umma launches of the run; the
# second forward starts at 28): raw page + source page as CSV
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
IDX=${1:-42}
TAGN=${2:-c64k3}
timeout 300 python tools/profile_target.py 2 128 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pair_umma -s $IDX -c 1 -o gpurun_out/src_$TAGN -f python tools/profile_target.py 2 128 > gpurun_out/ncu_src.log 2>&1
echo "ncu rc=$?"
ncu -i gpurun_out/src_$TAGN.ncu-rep --page raw --csv > gpurun_out/src_${TAGN}_raw.csv 2>/dev/null
ncu -i gpurun_out/src_$TAGN.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/src_${TAGN}_source.csv 2>/dev/null || ncu -i gpurun_out/src_$TAGN.ncu-rep --page source --csv > gpurun_out/src_${TAGN}_source.csv 2>/dev/null
ls -la gpurun_out/src_$TAGN*
rm -f gpurun_out/src_$TAGN.ncu-rep
