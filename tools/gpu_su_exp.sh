#!/bin/bash
# what-if timings of stream_umma_kernel (experiments build) + one ncu --set full capture of its five launches
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
X=$PWD/visual_onoma_to_wave_b200/libvo2w_x.so
run() { TAG="$1" timeout 300 python tools/launch_times.py 128 2>&1 | grep -E "^tag|^us:" | sed -E 's/ 2:[0-9]+//g'; }
run base
VO2W_LIB=$X run xbase
VO2W_LIB=$X VO2W_SU_DBG=1 run nostore
VO2W_LIB=$X VO2W_SU_DBG=2 run nomma
VO2W_LIB=$X VO2W_SU_DBG=3 run neither
for ng in 1 2 4 8; do VO2W_LIB=$X VO2W_SU_NG=$ng run ng$ng; done
timeout 300 python tools/profile_target.py 2 128 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stream_umma -s 5 -c 5 -o gpurun_out/su_prof -f python tools/profile_target.py 2 128 > gpurun_out/ncu_su.log 2>&1
echo "ncu rc=$?"
ncu -i gpurun_out/su_prof.ncu-rep --page raw --csv > gpurun_out/su_prof_raw.csv 2>/dev/null
ncu -i gpurun_out/su_prof.ncu-rep --page details --csv > gpurun_out/su_prof_details.csv 2>/dev/null
rm -f gpurun_out/su_prof.ncu-rep
