#!/bin/bash
# host-path tests, then the bench line without extras (value, ms per step, e2e)
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -q -m gpu -x -k "host or pipeline or int16 or vocoder or buffer" 2>&1 | tail -3
for i in 1 2; do
timeout 600 python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], d['clocks'])"
done
