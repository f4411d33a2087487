#!/bin/bash
# scratch experiment script (GPU box)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "tf32 or decoder or acoustic or postnet" 2>&1 | grep -E "AssertionError: |passed|failed|^FAILED|Error" | head -30 > gpurun_out/exp_tests.log
cat gpurun_out/exp_tests.log
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -1
timeout 300 python - > gpurun_out/exp_err.log 2>&1 <<'PY'
import torch, numpy as np
import visual_onoma_to_wave_b200 as v
from visual_onoma_to_wave_b200.synthetic import seeded_postnet_state_dict, seeded_decoder_state_dict, decoder_config
from oracle import postnet_oracle as po, decoder_oracle as do
g = np.load("tests/golden/postnet.npz")
pn = v.PostNet(); pn.load_state_dict(seeded_postnet_state_dict(4321)); pn.eval().cuda()
for name in "abc":
    x = torch.from_numpy(g[f"x_{name}"]).cuda(); ref = torch.from_numpy(g[f"y_{name}"])
    print("postnet golden", name, float((pn(x).cpu() - ref).abs().max()), "ref max", float(ref.abs().max()))
sd = seeded_postnet_state_dict(99); pn = v.PostNet(); pn.load_state_dict(sd); pn.eval().cuda()
x = torch.randn(64, 490, 80, generator=torch.Generator().manual_seed(5)) * 2 - 3
ref = po.postnet_forward(sd, x); got = pn(x.cuda()).cpu()
print("postnet cfg3", float((got - ref).abs().max()), "ref max", float(ref.abs().max()), "rms err", float((got-ref).pow(2).mean().sqrt()))
g = np.load("tests/golden/decoder.npz")
dec = v.Decoder(decoder_config()); dec.load_state_dict(seeded_decoder_state_dict(2468)); dec.eval().cuda()
y, _ = dec(torch.from_numpy(g["x_full"]).cuda(), torch.from_numpy(g["mask_full"]).cuda())
print("decoder golden full", float((y.cpu() - torch.from_numpy(g["y_full"])).abs().max()))
kw = dict(n_layers=6); sdd = seeded_decoder_state_dict(5, **kw)
dec = v.Decoder(decoder_config(**kw)); dec.load_state_dict(sdd); dec.eval().cuda()
lens = torch.tensor([490, 257, 64, 65, 1, 128, 300, 33]); L = 490
x = torch.randn(8, L, 256, generator=torch.Generator().manual_seed(3)); mask = torch.arange(L)[None, :] >= lens[:, None]
ref, _ = do.decoder_forward(sdd, x, mask); got, _ = dec(x.cuda(), mask.cuda())
print("decoder 6 layers ragged", float((got.cpu() - ref).abs().max()))
PY
cat gpurun_out/exp_err.log
