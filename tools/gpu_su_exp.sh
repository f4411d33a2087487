#!/bin/bash
# what-if timings of stream_umma_kernel (experiments build): single-layer launches only (conv_pre, ups0 .. ups3)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
X=$PWD/visual_onoma_to_wave_b200/libvo2w_x.so
run() { TAG="$1" timeout 300 python tools/launch_times.py 128 2>&1 | grep -E "^tag|^us:" | sed -E 's/ 2:[0-9]+//g'; }
export VO2W_LIB=$X
run xbase
VO2W_SU_DBG=1 run nostore
VO2W_SU_DBG=2 run nomma
VO2W_SU_DBG=3 run neither
VO2W_SU_SLOTS=4 run slots4
for ng in 1 2 4 8; do VO2W_SU_NG=$ng run ng$ng; done
VO2W_SU_PF=0 run pf0
VO2W_SU_PF=2 run pf2
run xbase2
