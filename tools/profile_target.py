"""Short, deterministic target for ncu: N forwards of Generator at T=172 (argv: N, B = 16, precision = bf16)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import visual_onoma_to_wave_b200 as v
from visual_onoma_to_wave_b200.synthetic import V1_CONFIG, seeded_state_dict, synthetic_mel

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
B = int(sys.argv[2]) if len(sys.argv) > 2 else 16
precision = sys.argv[3] if len(sys.argv) > 3 else "bf16"
dev = torch.device("cuda:0")
g = v.Generator(v.AttrDict(V1_CONFIG), precision=precision)
g.eval()
g.remove_weight_norm()
g.load_state_dict(seeded_state_dict(1234))
g.to(dev)
mel = synthetic_mel(B, 172, seed=0).to(dev)
for _ in range(n):
    wav = g(mel)
torch.cuda.synchronize()
print("ok", tuple(wav.shape), float(wav.abs().mean()), "launches/forward", g.last_launches)
