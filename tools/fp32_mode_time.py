"""Device time of Generator.forward in fp32 mode (max-abs-err <= 1e-4 gate) at configs[1]/[3] batch shapes."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import visual_onoma_to_wave_b200 as v
from visual_onoma_to_wave_b200.synthetic import V1_CONFIG, seeded_state_dict, synthetic_mel

g = v.Generator(v.AttrDict(V1_CONFIG), precision="fp32")
g.eval()
g.remove_weight_norm()
g.load_state_dict(seeded_state_dict(1234))
g.cuda()
for B in (16, 64):
    mel = synthetic_mel(B, 172, seed=1).cuda()
    for _ in range(2):
        g(mel)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(3):
        g(mel)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print(f"fp32 mode B={B} T=172: {ms:.2f} ms  {B * 172 * 256 / 22050 / ms * 1e3:.0f} audio-s/s  "
          f"{614105088 * B * 172 / ms / 1e9:.1f} TFLOP/s fp32-equivalent, launches {g.last_launches}")
