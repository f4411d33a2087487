"""Device time of whole Generator forwards at one batch size (CUDA events around runs of 8 forwards, no per-launch events)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import visual_onoma_to_wave_b200 as v
from visual_onoma_to_wave_b200.synthetic import V1_CONFIG, seeded_state_dict, synthetic_mel

B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
dev = torch.device("cuda:0")
g = v.Generator(v.AttrDict(V1_CONFIG), precision=sys.argv[2] if len(sys.argv) > 2 else "bf16")
g.eval()
g.remove_weight_norm()
g.load_state_dict(seeded_state_dict(1234))
g.to(dev)
mel = synthetic_mel(B, 172, seed=0).to(dev)
for _ in range(4):
    wav = g(mel)
torch.cuda.synchronize()
ms = []
for _ in range(7):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(8):
        wav = g(mel)
    b.record()
    torch.cuda.synchronize()
    ms.append(a.elapsed_time(b) / 8)
ms.sort()
print("tag", os.environ.get("TAG", ""), "B", B, "forward_ms min %.3f median %.3f max %.3f" % (ms[0], ms[len(ms) // 2], ms[-1]),
      "checksum %.6f" % float(wav.double().abs().mean()))
