#!/bin/bash
timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -3
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -1
python tools/launch_times.py 64 172 fp32 2>&1 | tail -2 | cut -c1-200
