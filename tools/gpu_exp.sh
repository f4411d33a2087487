#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -15
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -1
timeout 300 python - <<'PY'
import torch, time
import visual_onoma_to_wave_b200 as v
from visual_onoma_to_wave_b200.synthetic import V1_CONFIG, seeded_state_dict, synthetic_mel
from oracle import hifigan_oracle as ho
sd = seeded_state_dict(1234)
g = v.Generator(v.AttrDict(V1_CONFIG), precision="fp32"); g.eval(); g.remove_weight_norm(); g.load_state_dict(sd); g.cuda()
mel = synthetic_mel(4, 172, seed=0)
ref = ho.generator_forward(sd, mel)
wav = g(mel.cuda()).cpu()
print("fp32 mode B=4 T=172 max-abs-err", float((wav - ref).abs().max()), "ref max", float(ref.abs().max()))
ref64 = ho.generator_forward(sd, mel.double())
print("   oracle fp32 vs fp64", float((ref.double() - ref64).abs().max()), " gpu vs fp64", float((wav.double() - ref64).abs().max()))
for B in (16, 64):
    mel = synthetic_mel(B, 172, seed=1).cuda()
    for _ in range(2): g(mel)
    torch.cuda.synchronize(); e0 = torch.cuda.Event(True); e1 = torch.cuda.Event(True); e0.record()
    for _ in range(3): g(mel)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print(f"fp32 mode B={B} T=172: {ms:.2f} ms  {B*172*256/22050/ms*1e3:.0f} audio-s/s  {614105088*B*172/ms/1e9:.1f} TFLOP/s fp32-equivalent")
PY
