#!/bin/bash
mkdir -p gpurun_out
echo "--- n_t 128 for k=1"; PYTHONPATH=. python tools/dec_err.py 2>&1 | tail -5
echo "--- n_t 128 for all"; VO2W_T32_WIDE=1 PYTHONPATH=. python tools/dec_err.py 2>&1 | tail -5
python tools/bench_tail.py 2>/dev/null | grep "^{" | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print({k:round(v['ms'],2) for k,v in d['stages'].items()}, 'chain', round(d['chain_ms'],2))"
VO2W_T32_WIDE=1 python tools/bench_tail.py 2>/dev/null | grep "^{" | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('wide all', {k:round(v['ms'],2) for k,v in d['stages'].items()}, 'chain', round(d['chain_ms'],2))"
