"""Small, fast known-answer invocations of every hand-written kernel family, for compute-sanitizer (tools/sanitize.sh):
stream_umma_kernel, umma_conv_kernel, pair_umma_kernel (with and without conv_post on board), tf32_conv_kernel, the fp32
CUDA-core convolutions, LengthRegulator, and one tiny Generator forward per precision mode.  Results are checked against
torch so that a sanitizer run that silently changed behaviour is noticed.

    python tools/sanitize_target.py [family ...]      families: stream umma pair post tf32 simt lr gen
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402
import visual_onoma_to_wave_b200 as v  # noqa: E402
from visual_onoma_to_wave_b200 import _lib  # noqa: E402
from visual_onoma_to_wave_b200.synthetic import V1_CONFIG, seeded_state_dict, synthetic_mel  # noqa: E402

dev = torch.device("cuda:0")
lib = _lib.lib()
b16 = lambda t: t.to(torch.bfloat16).to(t.dtype)


def layer(umma, tr, Cin, Cout, k, ds, L, B):
    g = torch.Generator().manual_seed(Cin + k)
    w = torch.randn((Cin, Cout, k) if tr else (Cout, Cin, k), generator=g) / (Cin * k) ** 0.5
    b = torch.randn(Cout, generator=g) * 0.1
    x = torch.randn(B, Cin, L, generator=g)
    out = torch.full((B, Cout, L * (ds if tr else 1)), float("nan"), device=dev)
    xd = x.to(dev)
    _lib.check(lib.vo2w_debug_layer(tr, umma, w.data_ptr(), b.data_ptr(), Cin, Cout, k, ds, xd.data_ptr(), B, L, 0.1, None,
                                    out.data_ptr(), None))
    xa, wr = F.leaky_relu(x, 0.1), w
    if umma:
        xa, wr = b16(xa), b16(w)
    ref = (F.conv_transpose1d(xa.double(), wr.double(), b.double(), stride=ds, padding=(k - ds) // 2) if tr else
           F.conv1d(xa.double(), wr.double(), b.double(), padding=ds * (k - 1) // 2, dilation=ds))
    err = float((out.cpu().double() - ref).abs().max())
    assert err < 3e-5, err


def pair(C_, ks, ds, B, L, post=False):
    n = len(ks)
    g = torch.Generator().manual_seed(C_ + L)
    w1 = [(torch.randn(C_, C_, k, generator=g) / (C_ * k) ** 0.5).contiguous() for k in ks]
    w2 = [(torch.randn(C_, C_, k, generator=g) / (C_ * k) ** 0.5).contiguous() for k in ks]
    b1 = [torch.randn(C_, generator=g) * 0.1 for _ in ks]
    b2 = [torch.randn(C_, generator=g) * 0.1 for _ in ks]
    x = [torch.randn(B, C_, L, generator=g) for _ in ks]
    xd = [t.to(dev) for t in x]
    pa = lambda ts: (C.c_void_p * len(ts))(*[t.data_ptr() for t in ts])
    kk, dd = (C.c_int * n)(*ks), (C.c_int * n)(*ds)
    act = torch.full((B, C_, L), float("nan"), device=dev)
    _lib.check(lib.vo2w_debug_pair(n, pa(w1), pa(b1), pa(w2), pa(b2), C_, kk, dd, 2, pa(xd), B, L, 1.0 / n, 0.1 if not post else 0.01,
                                   None, act.data_ptr(), None))
    ref = 0
    for s in range(n):
        xa = b16(F.leaky_relu(x[s], 0.1)).double()
        t = F.conv1d(xa, b16(w1[s]).double(), b1[s].double(), padding=ds[s] * (ks[s] - 1) // 2, dilation=ds[s])
        t = b16(F.leaky_relu(t.float(), 0.1)).double()
        ref = ref + F.conv1d(t, b16(w2[s]).double(), b2[s].double(), padding=(ks[s] - 1) // 2) + torch.where(xa < 0, xa * 10.0, xa)
    aref = F.leaky_relu(ref / n, 0.01 if post else 0.1)
    got = act.cpu().double()
    assert int(((got - aref).abs() > aref.abs() * 2.0 ** -7 + 2e-2).sum()) == 0
    if post:
        pw = (torch.randn(1, 32, 7, generator=g) / 224 ** 0.5).contiguous()
        wav = torch.full((B, L), float("nan"), device=dev)
        i16 = torch.zeros((B, L), device=dev, dtype=torch.int16)
        _lib.check(lib.vo2w_debug_pair_post(n, pa(w1), pa(b1), pa(w2), pa(b2), kk, dd, pa(xd), B, L, 1.0 / n, 0.01, pw.data_ptr(), 0.05,
                                            wav.data_ptr(), i16.data_ptr(), 32768.0, None))
        r2 = torch.tanh(F.conv1d(got, pw.double(), torch.tensor([0.05]).double(), padding=3))[:, 0]
        assert float((wav.cpu().double() - r2).abs().max()) <= 5e-3


def tf32(B, L, Cin, Cout, k):
    g = torch.Generator().manual_seed(Cin + Cout)
    x = torch.randn(B, L, Cin, generator=g)
    w = torch.randn(Cout, Cin, k, generator=g) / (Cin * k) ** 0.5
    bias = torch.randn(Cout, generator=g)
    ref = F.conv1d(x.double().transpose(1, 2), w.double(), bias.double(), padding=(k - 1) // 2).transpose(1, 2)
    out = torch.full((B, L, Cout), float("nan"), device=dev)
    xd, wd, bd = x.to(dev), w.to(dev), bias.to(dev)
    scratch = torch.empty(2 * w.numel(), device=dev)
    _lib.check(lib.vo2w_debug_tf32_conv(xd.data_ptr(), wd.data_ptr(), bd.data_ptr(), None, out.data_ptr(), B, L, Cin, Cout, k, 0,
                                        scratch.data_ptr(), torch.cuda.current_stream(dev).cuda_stream))
    assert float((out.cpu().double() - ref).abs().max()) <= 4e-6 + 1e-8 * Cin * k


def gen():
    from oracle import hifigan_oracle as ho
    sd = seeded_state_dict(1234)
    mel = synthetic_mel(2, 12, seed=0)
    ref = ho.generator_forward(sd, mel)
    for prec in ("bf16", "fp32"):
        gg = v.Generator(v.AttrDict(V1_CONFIG), precision=prec)
        gg.eval()
        gg.remove_weight_norm()
        gg.load_state_dict(sd)
        gg.to(dev)
        wav = gg(mel.to(dev)).cpu()
        assert float((wav - ref).abs().max()) <= (1e-4 if prec == "fp32" else 3e-3)


def lr():
    x = torch.randn(4, 6, 256)
    d = torch.randint(0, 5, (4, 6))
    out, ml = v.LengthRegulator()(x.to(dev), d.to(dev), None)
    idx = torch.repeat_interleave(torch.arange(6).repeat(4), d.reshape(-1))
    assert int(ml.sum()) == idx.numel()


FAMILIES = {
    "stream": lambda: [layer(2, 1, 128, 64, 4, 2, 700, 2), layer(2, 1, 256, 128, 16, 8, 300, 2), layer(2, 0, 80, 512, 7, 1, 172, 3)],
    "umma": lambda: [layer(1, 0, 32, 1, 7, 1, 300, 2), layer(1, 1, 64, 32, 4, 2, 129, 1)],
    "pair": lambda: [pair(32, [3], [1], 2, 1200), pair(64, [7], [3], 2, 900), pair(128, [3, 7, 11], [5, 5, 5], 1, 700), pair(256, [3], [1], 2, 300)],
    "post": lambda: [pair(32, [3, 7, 11], [5, 5, 5], 2, 1300, post=True)],
    "tf32": lambda: [tf32(2, 300, 256, 768, 1), tf32(2, 131, 80, 512, 5)],
    "simt": lambda: [layer(0, 0, 32, 32, 3, 1, 200, 2), layer(0, 1, 64, 32, 4, 2, 129, 1)],
    "lr": lr,
    "gen": gen,
}

if __name__ == "__main__":
    names = sys.argv[1:] or list(FAMILIES)
    for nme in names:
        FAMILIES[nme]()
        torch.cuda.synchronize()
        print("ok", nme, flush=True)
