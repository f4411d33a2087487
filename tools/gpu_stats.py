"""Prints the statistics the shape-parity tests gate on (tests/test_gpu_shapes.py), so thresholds are set from
measurements: GPU vs the CPU restatement of its dataflow, per sample / per window / per stage."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import visual_onoma_to_wave_b200 as v  # noqa: E402
from visual_onoma_to_wave_b200.synthetic import V1_CONFIG, seeded_state_dict, synthetic_mel  # noqa: E402
from oracle import hifigan_oracle as ho  # noqa: E402
from oracle.emulate import generator_forward_emulated_v3, snr_db, ac_snr_db  # noqa: E402

dev = torch.device("cuda:0")
sd = seeded_state_dict(1234)
g = v.Generator(v.AttrDict(V1_CONFIG), precision="bf16")
g.eval(); g.remove_weight_norm(); g.load_state_dict(sd); g.to(dev)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
mel = synthetic_mel(B, 172, seed=0)
taps = {}
emu = generator_forward_emulated_v3(sd, mel, taps=taps)
ref = ho.generator_forward(sd, mel)
wav = g(mel.to(dev)).cpu()
for b in range(B):
    w, e, r = wav[b, 0].double(), emu[b, 0].double(), ref[b, 0].double()
    sigma = float((e - r).pow(2).mean().sqrt())
    d = w - e
    n = d.numel() // 128 * 128
    win = d[:n].reshape(-1, 128).pow(2).mean(1).sqrt()
    print(f"utt {b}: sigma_ref {sigma:.3e}  rms(gpu-emu)/sigma {float(d.pow(2).mean().sqrt()) / sigma:.3f}  max/sigma {float(d.abs().max()) / sigma:.2f}  "
          f"win max/sigma {float(win.max()) / sigma:.2f}  snr {snr_db(r, w):.2f} ac {ac_snr_db(r, w):.2f} emu snr {snr_db(r, e):.2f}")
for name in [f"ups.{i}" for i in range(4)] + [f"mrf_act.{i}" for i in range(4)]:
    if not g.has_tap(name):
        print(name, "no tap (fused)")
        continue
    got = g.debug_tap(name)
    exp = taps[name]
    got = got.reshape(exp.shape)
    rms = float(exp.double().pow(2).mean().sqrt())
    d = got.double() - exp.double()
    L = d.shape[-1] // 128 * 128
    win = d[..., :L].reshape(d.shape[0], d.shape[1], -1, 128).pow(2).mean(dim=(1, 3)).sqrt()
    print(f"{name}: rms {rms:.4f}  err rms/rms {float(d.pow(2).mean().sqrt()) / rms:.5f}  window max/rms {float(win.max()) / rms:.5f}  max abs {float(d.abs().max()):.4f}")
