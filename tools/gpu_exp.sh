#!/bin/bash
# scratch experiment script (GPU box)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "tf32 or decoder or acoustic or postnet" 2>&1 | grep -E "AssertionError: |passed|failed|^FAILED|Error" | head -30 > gpurun_out/exp_tests.log
cat gpurun_out/exp_tests.log
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -2
timeout 300 python - > gpurun_out/exp_tf32.log 2>&1 <<'PY'
import torch
from visual_onoma_to_wave_b200 import _lib
dev = torch.device("cuda:0")
lib = _lib.lib()
for (B, L, Cin, Cout, k) in ((64, 479, 256, 1024, 9), (64, 479, 1024, 256, 1), (64, 479, 256, 768, 1), (64, 479, 256, 256, 1), (64, 479, 512, 512, 5), (64, 479, 256, 80, 1)):
    x = torch.randn(B, L, Cin, device=dev); w = torch.randn(Cout, Cin, k, device=dev) * 0.05; b = torch.randn(Cout, device=dev)
    out = torch.empty(B, L, Cout, device=dev); scratch = torch.empty(2 * w.numel(), device=dev)
    st = torch.cuda.current_stream().cuda_stream
    def run(): _lib.check(lib.vo2w_debug_tf32_conv(x.data_ptr(), w.data_ptr(), b.data_ptr(), None, out.data_ptr(), B, L, Cin, Cout, k, 0, scratch.data_ptr(), st))
    for _ in range(2): run()
    torch.cuda.synchronize(); e0 = torch.cuda.Event(True); e1 = torch.cuda.Event(True); e0.record()
    for _ in range(5): run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    fl = 2 * B * L * Cin * Cout * k
    print(f"B{B} L{L} {Cin}->{Cout} k{k}: {ms*1e3:.0f} us (incl. pack)  {fl/ms/1e9:.1f} TFLOP/s fp32-equivalent, {3*fl/ms/1e9:.0f} TF32 TFLOP/s")
PY
cat gpurun_out/exp_tf32.log
python tools/bench_tail.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print({k:round(v['ms'],2) for k,v in d['stages'].items()}, 'chain', round(d['chain_ms'],2))"
