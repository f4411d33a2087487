#!/bin/bash
# scratch experiment script (GPU box)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "variance or decoder or postnet" 2>&1 | tail -30 > gpurun_out/exp_tests.log
cat gpurun_out/exp_tests.log
