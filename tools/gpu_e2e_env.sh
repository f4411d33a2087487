#!/bin/bash
# one GPU: legacy default stream against a created stream for the device-resident loop
cd "$(dirname "$0")/.."
B="python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline"
timeout 600 $B 2>/dev/null | python tools/show_bench.py plain
VO2W_BENCH_STREAM=side timeout 600 $B 2>/dev/null | python tools/show_bench.py side
timeout 600 $B 2>/dev/null | python tools/show_bench.py plain2
VO2W_BENCH_STREAM=side timeout 600 $B 2>/dev/null | python tools/show_bench.py side2
