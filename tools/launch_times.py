"""Per-launch device times of one Generator.forward (CUDA events around every launch, no profiler)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import visual_onoma_to_wave_b200 as v
from visual_onoma_to_wave_b200.synthetic import V1_CONFIG, seeded_state_dict, synthetic_mel

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
T = int(sys.argv[2]) if len(sys.argv) > 2 else 172
dev = torch.device("cuda:0")
PREC = sys.argv[3] if len(sys.argv) > 3 else "bf16"
g = v.Generator(v.AttrDict(V1_CONFIG), precision=PREC)
g.eval()
g.remove_weight_norm()
g.load_state_dict(seeded_state_dict(1234))
g.to(dev)
mel = synthetic_mel(B, T, seed=0).to(dev)
for _ in range(3):
    g(mel)
best = None
for _ in range(5):
    ms, kinds = g.launch_times(mel)
    if best is None or sum(ms) < sum(best):
        best = ms
print("tag", os.environ.get("TAG", ""), "B", B, "T", T, "total_us %.1f" % (sum(best) * 1e3),
      "pair_us %.1f" % (sum(m for m, k in zip(best, kinds) if k == 2) * 1e3),
      "single_us %.1f" % (sum(m for m, k in zip(best, kinds) if k == 1) * 1e3))
print("us:", " ".join("%d:%.0f" % (k, m * 1e3) for m, k in zip(best, kinds)))
