#!/bin/bash
# scratch A/B runs: per-launch times of one forward at B = 128 for the product build (libvo2w.so) and every variant
# build libvo2w_<tag>.so listed in $1 (comma separated), alternating so that clock drift shows.  $2 = 1: parity tests first.
cd "$(dirname "$0")/.."
run() { TAG="$1" timeout 300 python tools/launch_times.py 128 2>&1 | grep -E "^tag|^us:"; }
[ "$2" = 1 ] && timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_shapes.py -x -q -k "fused_resblock or generator or conv_post or host_pipeline or mrf or pair" 2>&1 | tail -3
for i in 1 2 3; do
  run new$i
  for t in ${1//,/ }; do
    L=$PWD/visual_onoma_to_wave_b200/libvo2w_$t.so
    [ -f $L ] && VO2W_LIB=$L run $t$i
  done
done
