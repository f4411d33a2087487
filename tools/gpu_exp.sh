#!/bin/bash
timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -2
python tools/fp32_mode_time.py 2>&1 | tail -1
python tools/bench_tail.py 2>/dev/null | grep "^{" | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print({k:round(v['ms'],2) for k,v in d['stages'].items()}, 'chain', round(d['chain_ms'],2))"
