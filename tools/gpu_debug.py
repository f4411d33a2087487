"""First-contact diagnostics on a B200: every layer class through both kernel families against
torch, then the whole Generator against the oracle, printing errors instead of asserting.
Run:  python tools/gpu_debug.py > gpurun_out/debug.log 2>&1
"""
import ctypes as C
import os
import sys
import time
import traceback

os.environ.setdefault("VO2W_DEBUG_TAPS", "1")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch
import torch.nn.functional as F

from visual_onoma_to_wave_b200 import _lib, AttrDict, Generator, LengthRegulator
from visual_onoma_to_wave_b200.synthetic import V1_CONFIG, seeded_state_dict, synthetic_mel
from oracle import hifigan_oracle as ho
from oracle import length_regulator_oracle as lro
from oracle.emulate import generator_forward_emulated, snr_db

dev = torch.device("cuda:0")
lib = _lib.lib()


def layer_probe(transposed, umma, w, b, x, ds, slope=1.0, res=None):
    Cin = x.shape[1]
    Cout = w.shape[1] if transposed else w.shape[0]
    k = w.shape[2]
    B, _, L = x.shape
    Lo = L * (ds if transposed else 1)
    out = torch.full((B, Cout, Lo), float("nan"), device=dev)
    wc, bc = w.contiguous().cpu(), b.contiguous().cpu()
    xd = x.contiguous().to(dev)
    rd = res.contiguous().to(dev) if res is not None else None
    rc = lib.vo2w_debug_layer(int(transposed), int(umma), wc.data_ptr(), bc.data_ptr(), Cin, Cout, k, ds, xd.data_ptr(),
                              B, L, float(slope), rd.data_ptr() if rd is not None else None, out.data_ptr(), None)
    if rc != 0:
        raise RuntimeError(lib.vo2w_last_error().decode())
    torch.cuda.synchronize()
    return out.cpu()


def ref_layer(transposed, w, b, x, ds, slope, res, bf16):
    xa = F.leaky_relu(x, slope) if slope != 1.0 else x
    if bf16:
        xa, w = xa.to(torch.bfloat16).float(), w.to(torch.bfloat16).float()
    xa, w, b = xa.double(), w.double(), b.double()
    if transposed:
        y = F.conv_transpose1d(xa, w, b, stride=ds, padding=ds // 2)
    else:
        y = F.conv1d(xa, w, b, padding=ds * (w.shape[2] - 1) // 2, dilation=ds)
    if res is not None:
        y = y + res.double()
    return y.float()


def main():
    sections = set(sys.argv[1:]) or {"simt", "umma", "lr", "gen_fp32", "gen_bf16"}
    print("sections", sorted(sections))
    print(torch.cuda.get_device_name(0), "lib version", lib.vo2w_version(), "device_check", lib.vo2w_device_check())
    g = torch.Generator().manual_seed(0)
    cases = [  # transposed, Cin, Cout, k, d_or_s, L, B, slope, res
        (0, 32, 32, 3, 1, 200, 2, 0.1, True), (0, 32, 32, 11, 5, 300, 1, 0.1, False), (0, 64, 64, 7, 3, 130, 2, 0.1, True),
        (0, 128, 128, 3, 3, 256, 1, 0.1, True), (0, 256, 256, 11, 5, 129, 1, 0.1, True), (0, 256, 256, 3, 1, 64, 2, 1.0, False),
        (0, 80, 512, 7, 1, 50, 2, 1.0, False), (0, 32, 1, 7, 1, 300, 2, 0.01, False),
        (1, 512, 256, 16, 8, 40, 1, 0.1, False), (1, 256, 128, 16, 8, 140, 1, 0.1, False),
        (1, 128, 64, 4, 2, 300, 2, 0.1, False), (1, 64, 32, 4, 2, 129, 1, 0.1, False),
    ]
    for (tr, Cin, Cout, k, ds, L, B, slope, use_res) in cases:
        w = torch.randn((Cin, Cout, k) if tr else (Cout, Cin, k), generator=g) / (Cin * k) ** 0.5
        b = torch.randn(Cout, generator=g) * 0.1
        x = torch.randn(B, Cin, L, generator=g)
        res = torch.randn(B, Cout, L, generator=g) if use_res else None
        for umma in (0, 1):
            if ("umma" if umma else "simt") not in sections:
                continue
            tag = f"{'convT' if tr else 'conv '} Cin={Cin} Cout={Cout} k={k} ds={ds} L={L} B={B} {'umma' if umma else 'simt'}"
            try:
                t0 = time.time()
                got = layer_probe(tr, umma, w, b, x, ds, slope, res)
                ref = ref_layer(tr, w, b, x, ds, slope, res, bf16=bool(umma))
                err = (got - ref).abs()
                nan = int(torch.isnan(got).sum())
                print(f"{tag}: max_err {float(err.nan_to_num(9e9).max()):.3e} mean_err {float(err.nan_to_num(0).mean()):.3e} "
                      f"nan {nan} ref_absmax {float(ref.abs().max()):.3f} ({time.time() - t0:.2f}s)")
                if float(err.nan_to_num(9e9).max()) > 1e-3:
                    bad = (err.nan_to_num(9e9) > 1e-3)
                    idx = bad.nonzero()[:5].tolist()
                    print("   first bad idx", idx, "got", [float(got[tuple(i)]) for i in idx], "ref", [float(ref[tuple(i)]) for i in idx])
                    print("   bad per batch", bad.sum((1, 2)).tolist(), "bad rows(l) sample", sorted(set(int(i[2]) for i in bad.nonzero()[:2000].tolist()))[:24],
                          "bad ch sample", sorted(set(int(i[1]) for i in bad.nonzero()[:2000].tolist()))[:24])
            except Exception as e:
                print(f"{tag}: EXCEPTION {e}")
                traceback.print_exc()

    # ---- LengthRegulator ----
    try:
        if "lr" not in sections:
            raise KeyError("skipped")
        lr = LengthRegulator()
        x = torch.randn(64, 32, 256, generator=g)
        d = torch.randint(0, 25, (64, 32), generator=g)
        out, ml = lr(x.to(dev), d.to(dev), None)
        eo, eml = lro.length_regulate(x.numpy(), d.numpy(), None)
        print("LR cfg3 exact:", bool((out.cpu().numpy() == eo).all()), bool((ml.cpu().numpy() == eml).all()), tuple(out.shape))
    except Exception as e:
        print("LR EXCEPTION", e)
        traceback.print_exc()

    # ---- whole Generator ----
    h = AttrDict(V1_CONFIG)
    sd = seeded_state_dict(1234)
    mel = synthetic_mel(2, 40, seed=0)
    taps = {}
    ref = ho.generator_forward(sd, mel, taps=taps)
    for prec in ("fp32", "bf16"):
        if "gen_" + prec not in sections:
            continue
        try:
            gen = Generator(h, precision=prec)
            gen.eval()
            gen.remove_weight_norm()
            gen.load_state_dict(sd)
            gen.to(dev)
            t0 = time.time()
            wav = gen(mel.to(dev))
            torch.cuda.synchronize()
            t1 = time.time()
            wav = wav.cpu()
            print(f"generator {prec}: max_err {float((wav - ref).abs().max()):.3e} snr {snr_db(ref, wav):.1f} dB launches {gen.last_launches} first-call {t1 - t0:.3f}s nan {int(torch.isnan(wav).sum())}")
            for name in ["conv_pre"] + [x for i in range(4) for x in (f"ups.{i}", f"mrf.{i}")]:
                try:
                    t = gen.debug_tap(name).reshape(taps[name].shape)
                    e = (t - taps[name]).abs()
                    print(f"   tap {name}: max_err {float(e.max()):.3e} ref_absmax {float(taps[name].abs().max()):.3f} snr {snr_db(taps[name], t):.1f}")
                except Exception as e:
                    print(f"   tap {name}: {e}")
            if prec == "bf16":
                emu = generator_forward_emulated(sd, mel, mode="bf16", acc_dtype=torch.float32)
                print(f"   vs emulation: max_err {float((wav - emu).abs().max()):.3e} snr {snr_db(emu, wav):.1f} dB")
            # timing at config-2 size
            melb = synthetic_mel(16, 172, seed=1).to(dev)
            for _ in range(3):
                y = gen(melb)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n = 3 if prec == "fp32" else 10
            e0.record()
            for _ in range(n):
                y = gen(melb)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / n
            print(f"   cfg2 B=16 T=172: {ms:.3f} ms/forward -> {16 * 172 * 256 / 22050 / (ms / 1e3):.0f} audio-s/s, "
                  f"{614105088 * 16 * 172 / (ms / 1e3) / 1e12:.1f} TFLOP/s")
            del gen
        except Exception as e:
            print(f"generator {prec}: EXCEPTION {e}")
            traceback.print_exc()


if __name__ == "__main__":
    main()
