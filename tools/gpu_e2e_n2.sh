#!/bin/bash
# two ranks on two GPUs: e2e / value with the NCCL process group (as bench.py runs) and with VO2W_BENCH_PG=gloo
cd "$(dirname "$0")/.."
for pg in nccl gloo nccl gloo; do
VO2W_BENCH_PG=$pg timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 3 --warmup 3 --no-extras --no-cpu-baseline 2>/dev/null | python tools/show_bench.py $pg
done
