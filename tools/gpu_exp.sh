#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "tf32 or decoder or acoustic or postnet" 2>&1 | grep -E "AssertionError: |passed|failed|^FAILED|Error" | head -30
PYTHONPATH=. python tools/dec_err.py 2>&1 | tail -5
python tools/bench_tail.py 2>/dev/null | grep "^{" | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print({k:round(v['ms'],2) for k,v in d['stages'].items()}, 'chain', round(d['chain_ms'],2))"
