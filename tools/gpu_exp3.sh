#!/bin/bash
# scratch A/B of whole-forward times: environment variants of the product build, alternating
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_shapes.py -x -q 2>&1 | tail -3
for i in 1 2 3; do
  TAG=pdl$i timeout 300 python tools/forward_time.py 128 2>&1 | grep "^tag"
  TAG=nopdl$i VO2W_PDL=0 timeout 300 python tools/forward_time.py 128 2>&1 | grep "^tag"
done
TAG=pdl_b16 timeout 300 python tools/forward_time.py 16 2>&1 | grep "^tag"
TAG=nopdl_b16 VO2W_PDL=0 timeout 300 python tools/forward_time.py 16 2>&1 | grep "^tag"
