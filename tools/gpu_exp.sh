#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out



for sb in 64 128 256; do timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --sub-batch $sb 2>&1 | grep "^{" | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('sub', d['config']['sub_batch'], 'value', d['value'], 'e2e', d['e2e']['value'], 'roof', d['roofline']['frac'], d['clocks'])"; done
