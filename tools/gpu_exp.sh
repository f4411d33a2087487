#!/bin/bash
# scratch experiment script (GPU box)
mkdir -p gpurun_out
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
# fp64 arbiter: how far is the fp32 CPU oracle itself from fp64?
sd64 = {k: (t.double() if t.is_floating_point() else t) for k, t in sd.items()}
ref64 = po.postnet_forward(sd64, x.double())
print("  oracle fp32 vs fp64", float((ref.double() - ref64).abs().max()), " gpu vs fp64", float((got.double() - ref64).abs().max()))
kw = dict(n_layers=6); sdd = seeded_decoder_state_dict(5, **kw)
dec = v.Decoder(decoder_config(**kw)); dec.load_state_dict(sdd); dec.eval().cuda()
lens = torch.tensor([490, 257, 64, 65, 1, 128, 300, 33]); L = 490
x = torch.randn(8, L, 256, generator=torch.Generator().manual_seed(3)); mask = torch.arange(L)[None, :] >= lens[:, None]
ref, _ = do.decoder_forward(sdd, x, mask); got, _ = dec(x.cuda(), mask.cuda())
sd64 = {k: t.double() for k, t in sdd.items()}
ref64, _ = do.decoder_forward(sd64, x.double(), mask)
print("decoder 6 layers ragged", float((got.cpu() - ref).abs().max()), "oracle fp32 vs fp64", float((ref.double() - ref64).abs().max()), "gpu vs fp64", float((got.cpu().double() - ref64).abs().max()))
PY
cat gpurun_out/exp_err.log
