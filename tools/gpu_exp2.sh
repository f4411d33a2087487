#!/bin/bash
cd "$(dirname "$0")/.."
X=$PWD/visual_onoma_to_wave_b200/libvo2w_x.so
run() { TAG="$1" timeout 300 python tools/launch_times.py 128 2>&1 | grep -E "^tag|^us:"; }
VO2W_LIB=$X timeout 900 python -m pytest tests/test_gpu_parity.py -q -k "fused_resblock or generator or conv_post or host_pipeline" 2>&1 | tail -2
VO2W_LIB=$X run e1init
VO2W_LIB=$X VO2W_PAIR_E1INIT=0 run off
VO2W_LIB=$X run e1init2
VO2W_LIB=$X VO2W_PAIR_E1INIT=0 run off2
