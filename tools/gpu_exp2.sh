#!/bin/bash
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_gpu_parity.py -q -k "fused_resblock or generator or conv_post or host_pipeline" 2>&1 | tail -2
run() { TAG="$1" timeout 300 python tools/launch_times.py 128 2>&1 | grep -E "^tag|^us:"; }
run nodiv
run nodiv2
