#!/bin/bash
cd "$(dirname "$0")/.."
X=$PWD/visual_onoma_to_wave_b200/libvo2w_x.so   # build with VO2W_NVCC_EXTRA="-DVO2W_EXPERIMENTS -DVO2W_PAIR_PROFILE"
VO2W_LIB=$X VO2W_PAIR_PROF=1 VO2W_PAIR_TRACE=1 timeout 300 python tools/profile_target.py 1 128 2>&1 | grep -A10 "pair prof\] C=64 segs=1 k0=3\|pair prof\] C=32 segs=1 k0=3" | cut -c1-330
