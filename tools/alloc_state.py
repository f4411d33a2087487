"""Does the speed of the memory-bound launches depend on where the workspace lands?  Re-allocates the Generator workspace
several times in ONE process (different dummy allocations in front of it) and prints per-launch times of each placement."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import visual_onoma_to_wave_b200 as v
from visual_onoma_to_wave_b200.synthetic import V1_CONFIG, seeded_state_dict, synthetic_mel

dev = torch.device("cuda:0")
g = v.Generator(v.AttrDict(V1_CONFIG), precision="bf16")
g.eval()
g.remove_weight_norm()
g.load_state_dict(seeded_state_dict(1234))
g.to(dev)
mel = synthetic_mel(128, 172, seed=0).to(dev)
dummies = []
for trial in range(8):
    g._ws.clear()
    torch.cuda.empty_cache()
    if trial % 2 == 1:
        dummies.append(torch.empty((trial * 37 + 11) << 20, dtype=torch.uint8, device=dev))   # shifts the next cudaMalloc
    for _ in range(3):
        g(mel)
    best = None
    for _ in range(4):
        ms, kinds = g.launch_times(mel)
        if best is None or sum(ms) < sum(best):
            best = ms
    ws = next(iter(g._ws.values()))
    us = [int(m * 1e3) for m in best]
    print("trial", trial, "ws ptr %#x" % ws.data_ptr(), "total %.0f" % (sum(best) * 1e3), "ups1 %d ups2 %d ups3 %d" % (us[10], us[18], us[26]),
          "C128k3 %d %d" % (us[11], us[12]), "C64k3 %d %d" % (us[19], us[20]), "C32 %s" % " ".join(map(str, us[27:34])))
