#!/bin/bash
# scratch A/B runs: per-launch times of one forward at B = 128, product build (libvo2w.so) against a build of the
# previous commit (libvo2w_prev.so, when present), alternating so that clock drift shows
cd "$(dirname "$0")/.."
run() { TAG="$1" timeout 300 python tools/launch_times.py 128 2>&1 | grep -E "^tag|^us:"; }
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_shapes.py -x -q -k "fused_resblock or generator or conv_post or host_pipeline or mrf or pair" 2>&1 | tail -3
P=$PWD/visual_onoma_to_wave_b200/libvo2w_prev.so
for i in 1 2 3; do
  run new$i
  [ -f $P ] && VO2W_LIB=$P run prev$i
done
