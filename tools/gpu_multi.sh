#!/bin/bash
# multi-GPU checks (gpurun --gpus N): two devices from one process, then the driver's launch of bench.py at N ranks
cd "$(dirname "$0")/.."
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 600 python -m pytest tests/test_gpu_shapes.py -q -k "two_gpus" > gpurun_out/t_2gpu.log 2>&1; echo "two-gpu test rc=$?"; tail -n 3 gpurun_out/t_2gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_n$N.log 2> gpurun_out/bench_n$N.err; echo "bench N=$N rc=$?"
tail -n 1 gpurun_out/bench_n$N.log | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print({k:d[k] for k in ('value','n_gpus','ms_per_step','scaling')}, 'e2e', d['e2e']['value'], 'strong', d['extra'].get('strong_scaling'))
"
tail -n 3 gpurun_out/bench_n$N.err
