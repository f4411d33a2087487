#!/bin/bash
timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -3
echo "--- current"; PYTHONPATH=. python tools/dec_err.py 2>&1 | tail -5
python tools/fp32_mode_time.py 2>&1 | tail -2
