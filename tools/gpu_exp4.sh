#!/bin/bash
# scratch: re-tune the launch parameters of pair_umma_kernel with the experiments build (environment switches)
cd "$(dirname "$0")/.."
X=$PWD/visual_onoma_to_wave_b200/libvo2w_x.so
run() { TAG="$1" timeout 300 python tools/launch_times.py 128 2>&1 | grep -E "^tag|^us:"; }
export VO2W_LIB=$X
for i in 1 2; do
run base$i
VO2W_PAIR_SPLIT=0 run split0_$i
VO2W_PAIR_XABUFS=2 run xabufs2_$i
VO2W_PAIR_PF=1 run pf1_$i
VO2W_PAIR_RESIDENT=0 run resident0_$i
done
run base3
