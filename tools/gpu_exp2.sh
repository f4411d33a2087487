#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -q -x -k "fused_resblock or conv_post_on_board or generator or host_pipeline" > gpurun_out/t_k.log 2>&1; echo "kernel tests rc=$?"; tail -n 4 gpurun_out/t_k.log
X=$PWD/visual_onoma_to_wave_b200/libvo2w_x.so
run() { TAG="$1" timeout 300 python tools/launch_times.py 128 2>&1 | grep -E "^tag|^us:"; }
VO2W_LIB=$X run rinit
VO2W_LIB=$X VO2W_PAIR_RINIT=0 run no_rinit
VO2W_LIB=$X run rinit2
VO2W_LIB=$X VO2W_PAIR_RINIT=0 run no_rinit2
