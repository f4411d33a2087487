#!/bin/bash
cd "$(dirname "$0")/.."
X=$PWD/visual_onoma_to_wave_b200/libvo2w_x.so
run() { TAG="$1" timeout 300 python tools/launch_times.py 128 2>&1 | grep -E "^tag|^us:"; }
VO2W_LIB=$X timeout 600 python -m pytest tests/test_gpu_parity.py -q -x -k "fused_resblock or generator" 2>&1 | tail -2
run base
VO2W_LIB=$X run two_in_flight
run base2
VO2W_LIB=$X run two_in_flight2
