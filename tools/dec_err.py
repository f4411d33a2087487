import torch, numpy as np
import visual_onoma_to_wave_b200 as v
from visual_onoma_to_wave_b200.synthetic import seeded_decoder_state_dict, decoder_config
from oracle import decoder_oracle as do
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
for B, L in ((64, 479), (16, 479), (64, 172)):
    x = torch.randn(B, L, 256, device="cuda"); m = torch.zeros(B, L, dtype=torch.bool, device="cuda")
    for _ in range(2): dec(x, m)
    torch.cuda.synchronize(); e0 = torch.cuda.Event(True); e1 = torch.cuda.Event(True)
    e0.record()
    for _ in range(5): dec(x, m)
    e1.record(); torch.cuda.synchronize()
    print(f"decoder B={B} L={L}: {e0.elapsed_time(e1) / 5:.2f} ms")
