#!/bin/bash
# scratch A/B of whole-forward times (sustained): product build against libvo2w_prev.so, alternating
cd "$(dirname "$0")/.."
P=$PWD/visual_onoma_to_wave_b200/libvo2w_prev.so
for i in 1 2 3 4; do
  TAG=new$i timeout 300 python tools/forward_time.py 128 2>&1 | grep "^tag"
  TAG=prev$i VO2W_LIB=$P timeout 300 python tools/forward_time.py 128 2>&1 | grep "^tag"
done
