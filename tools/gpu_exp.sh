#!/bin/bash
# scratch experiment script (GPU box)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "tf32 or decoder or acoustic or postnet or variance" 2>&1 | grep -E "AssertionError: |passed|failed|^FAILED|Error" | head -30 > gpurun_out/exp_tests.log
cat gpurun_out/exp_tests.log
python tools/bench_tail.py > gpurun_out/tail.json 2> gpurun_out/tail.err; tail -3 gpurun_out/tail.err; cat gpurun_out/tail.json
