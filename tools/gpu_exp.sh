#!/bin/bash
# scratch experiment script (GPU box)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "decoder or postnet" 2>&1 | tail -15 > gpurun_out/exp_tests.log
python - > gpurun_out/exp_dec.log 2>&1 <<'PY'
import torch, time
import visual_onoma_to_wave_b200 as v
from visual_onoma_to_wave_b200.synthetic import seeded_decoder_state_dict, decoder_config
dec = v.Decoder(decoder_config()); dec.load_state_dict(seeded_decoder_state_dict(1)); dec.eval().cuda()
for B, L in ((64, 490), (16, 490)):
    x = torch.randn(B, L, 256, device="cuda"); m = torch.zeros(B, L, dtype=torch.bool, device="cuda")
    for _ in range(2): dec(x, m)
    torch.cuda.synchronize(); e0 = torch.cuda.Event(True); e1 = torch.cuda.Event(True)
    e0.record()
    for _ in range(5): dec(x, m)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    fl = B * L * 6 * 2 * (4 * 256 * 256 + 2 * 2 * L * 128 + 1024 * 256 * 9 + 1024 * 256)
    print(f"decoder B={B} L={L}: {ms:.2f} ms  {fl / ms / 1e9:.1f} TFLOP/s  {B * L / ms * 1e3:.0f} frames/s")
PY
cat gpurun_out/exp_tests.log gpurun_out/exp_dec.log
