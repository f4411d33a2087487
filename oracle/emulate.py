"""Operand-rounding emulation of the CUDA bf16 pipeline (TEST INFRASTRUCTURE).

Restates what the sm_100a kernels compute, on CPU, so GPU results can be compared
layer by layer far below the end-to-end tolerance: conv operands (activations after
leaky-relu, and weights) rounded to bf16, products accumulated in fp32 (here fp64 then
cast), bias / residual / MRF average in fp32.  Also used to pick which tensors may be
stored in bf16 without breaking the 40 dB gate (numerics section).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from .hifigan_oracle import V1_CONFIG, LRELU_SLOPE, fold_weight_norm, get_padding


def _r(x, mode):
    if mode == "bf16":
        return x.to(torch.bfloat16).to(x.dtype)
    if mode == "tf32":
        # round-to-nearest-even to 10 mantissa bits
        i = x.float().view(torch.int32)
        i = (i + 0x0FFF + ((i >> 13) & 1)) & ~0x1FFF
        return i.view(torch.float32).to(x.dtype)
    return x


def generator_forward_emulated(sd, mel, cfg=V1_CONFIG, mode="bf16", residual="fp32",
                               acc_dtype=torch.float64, taps=None):
    """Same dataflow as oracle.generator_forward with operand rounding ``mode``.

    residual: "fp32" keeps the residual stream x in fp32 between conv pairs,
              "bf16" rounds it after every pair (what a bf16 residual buffer would do).
    """
    sd = fold_weight_norm(sd)
    W = {k: (_r(v.float(), mode) if k.endswith("weight") else v.float()).to(acc_dtype)
         for k, v in sd.items()}
    nk = len(cfg["resblock_kernel_sizes"])
    f32 = lambda t: t.float().to(acc_dtype)      # fp32 storage of an fp32-accumulated value

    def conv(x, name, **kw):
        return f32(F.conv1d(_r(x, mode), W[name + ".weight"], W[name + ".bias"], **kw))

    with torch.no_grad():
        x = conv(mel.to(acc_dtype), "conv_pre", padding=3)
        if taps is not None:
            taps["conv_pre"] = x.float()
        for i, (u, k) in enumerate(zip(cfg["upsample_rates"], cfg["upsample_kernel_sizes"])):
            x = F.leaky_relu(x, LRELU_SLOPE)
            x = f32(F.conv_transpose1d(_r(x, mode), W[f"ups.{i}.weight"], W[f"ups.{i}.bias"],
                                       stride=u, padding=(k - u) // 2))
            if taps is not None:
                taps[f"ups.{i}"] = x.float()
            xs = None
            for j, (rk, rd) in enumerate(zip(cfg["resblock_kernel_sizes"],
                                             cfg["resblock_dilation_sizes"])):
                p = f"resblocks.{i * nk + j}"
                r = x
                for m, d in enumerate(rd):
                    xt = conv(F.leaky_relu(r, LRELU_SLOPE), f"{p}.convs1.{m}",
                              padding=get_padding(rk, d), dilation=d)
                    xt = conv(F.leaky_relu(xt, LRELU_SLOPE), f"{p}.convs2.{m}",
                              padding=get_padding(rk, 1))
                    r = f32(xt + r)
                    if residual == "bf16" and m < 2:
                        r = _r(r, "bf16")
                if taps is not None:
                    taps[p] = r.float()
                xs = r if xs is None else f32(xs + r)
            x = f32(xs / nk)
            if taps is not None:
                taps[f"mrf.{i}"] = x.float()
        x = F.leaky_relu(x)
        x = conv(x, "conv_post", padding=3)
        x = torch.tanh(x)
    return x.float()


def snr_db(ref: torch.Tensor, got: torch.Tensor) -> float:
    """Total-power SNR, 10 log10(sum ref^2 / sum err^2) (hard parts)."""
    ref = ref.double()
    err = got.double() - ref
    return float(10 * torch.log10(ref.pow(2).sum() / err.pow(2).sum().clamp_min(1e-300)))


def ac_snr_db(ref: torch.Tensor, got: torch.Tensor) -> float:
    """SNR against the AC power of the reference alone, 10 log10(sum (ref - mean ref)^2 / sum err^2): the error
    is taken as it is (its own mean counts as error), only the reference's DC term is denied to the numerator."""
    ref = ref.double()
    err = got.double() - ref
    return float(10 * torch.log10((ref - ref.mean()).pow(2).sum() / err.pow(2).sum().clamp_min(1e-300)))


def generator_forward_emulated_v2(sd, mel, cfg=V1_CONFIG, acc_dtype=torch.float32, taps=None):
    """Dataflow of the pair-fused bf16 pipeline (DESIGN.md "Numerics"): every activation tensor
    that reaches HBM is raw bf16; leaky-relu is applied to the bf16 value and rounded to bf16 again
    for the MMA operand; conv1->conv2 intermediate is bf16(lrelu(fp32)); residual adds, bias and the
    MRF sum are fp32; the MRF average is stored as raw bf16."""
    sd = fold_weight_norm(sd)
    W = {k: (_r(v.float(), "bf16") if k.endswith("weight") else v.float()).to(acc_dtype) for k, v in sd.items()}
    nk = len(cfg["resblock_kernel_sizes"])
    f32 = lambda t: t.float().to(acc_dtype)
    b16 = lambda t: _r(t, "bf16")
    act = lambda t, s: b16(F.leaky_relu(t, s))          # bf16 in, bf16 out

    with torch.no_grad():
        x = b16(f32(F.conv1d(b16(mel.to(acc_dtype)), W["conv_pre.weight"], W["conv_pre.bias"], padding=3)))
        if taps is not None:
            taps["conv_pre"] = x.float()
        for i, (u, k) in enumerate(zip(cfg["upsample_rates"], cfg["upsample_kernel_sizes"])):
            x = b16(f32(F.conv_transpose1d(act(x, LRELU_SLOPE), W[f"ups.{i}.weight"], W[f"ups.{i}.bias"], stride=u,
                                           padding=(k - u) // 2)))
            if taps is not None:
                taps[f"ups.{i}"] = x.float()
            xs = None
            for j, (rk, rd) in enumerate(zip(cfg["resblock_kernel_sizes"], cfg["resblock_dilation_sizes"])):
                p = f"resblocks.{i * nk + j}"
                r = x
                for m, d in enumerate(rd):
                    t = f32(F.conv1d(act(r, LRELU_SLOPE), W[f"{p}.convs1.{m}.weight"], W[f"{p}.convs1.{m}.bias"],
                                     padding=get_padding(rk, d), dilation=d))
                    t = b16(F.leaky_relu(t, LRELU_SLOPE))
                    y = f32(F.conv1d(t, W[f"{p}.convs2.{m}.weight"], W[f"{p}.convs2.{m}.bias"], padding=get_padding(rk, 1)))
                    r = f32(y + r)
                    if m < 2:
                        r = b16(r)
                xs = r if xs is None else f32(xs + r)
            x = b16(f32(xs / nk))
            if taps is not None:
                taps[f"mrf.{i}"] = x.float()
        x = f32(F.conv1d(act(x, 0.01), W["conv_post.weight"], W["conv_post.bias"], padding=3))
        x = torch.tanh(x)
    return x.float()


def generator_forward_emulated_v3(sd, mel, cfg=V1_CONFIG, acc_dtype=torch.float32, taps=None, fused_post=True):
    """Dataflow of the single-stream bf16 pipeline: the ONLY tensor a layer leaves in HBM is
    s = bf16(leaky_relu(y_fp32, 0.1)), the next convolution's MMA operand.  A consumer that needs the
    value itself (the residual add of ResBlock.forward, models.py:102) recovers it exactly-invertibly
    from the stream: x = s if s >= 0 else 10 * s (an exact fp32 product of a bf16 number).  Compared with
    v2 the operand is rounded once instead of twice and half of the HBM traffic disappears.

    fused_post (the default schedule of the library): conv_post + tanh ride the last MRF launch, which holds
    leaky_relu(xs / num_kernels, 0.01) in fp32 registers: conv_post then runs in fp32 on the unrounded tensor with fp32
    weights (csrc/pair_umma.cu, POST) and "mrf_act.<last>" never exists.  fused_post=False restates the separate
    conv_post launch (bf16 stream in, bf16 weights; VO2W_FUSE_POST=0)."""
    sd_f32 = {k: v.float().to(acc_dtype) for k, v in fold_weight_norm(sd).items()}
    sd = fold_weight_norm(sd)
    W = {k: (_r(v.float(), "bf16") if k.endswith("weight") else v.float()).to(acc_dtype) for k, v in sd.items()}
    nk = len(cfg["resblock_kernel_sizes"])
    f32 = lambda t: t.float().to(acc_dtype)
    b16 = lambda t: _r(t, "bf16")
    stream = lambda t, s=LRELU_SLOPE: b16(F.leaky_relu(f32(t), s))         # fp32 value -> HBM stream
    inv = lambda s: f32(torch.where(s < 0, s * 10.0, s))                   # stream -> value (slope 0.1 only)

    with torch.no_grad():
        s = stream(F.conv1d(b16(mel.to(acc_dtype)), W["conv_pre.weight"], W["conv_pre.bias"], padding=3))
        for i, (u, k) in enumerate(zip(cfg["upsample_rates"], cfg["upsample_kernel_sizes"])):
            s = stream(F.conv_transpose1d(s, W[f"ups.{i}.weight"], W[f"ups.{i}.bias"], stride=u, padding=(k - u) // 2))
            if taps is not None:
                taps[f"ups.{i}"] = inv(s).float()
            xs = None
            for j, (rk, rd) in enumerate(zip(cfg["resblock_kernel_sizes"], cfg["resblock_dilation_sizes"])):
                p = f"resblocks.{i * nk + j}"
                sr = s
                for m, d in enumerate(rd):
                    t = f32(F.conv1d(sr, W[f"{p}.convs1.{m}.weight"], W[f"{p}.convs1.{m}.bias"],
                                     padding=get_padding(rk, d), dilation=d))
                    t = b16(F.leaky_relu(t, LRELU_SLOPE))
                    y = f32(F.conv1d(t, W[f"{p}.convs2.{m}.weight"], W[f"{p}.convs2.{m}.bias"], padding=get_padding(rk, 1)))
                    r = f32(y + inv(sr))
                    if m < 2:
                        sr = stream(r)
                xs = r if xs is None else f32(xs + r)
            last = i + 1 == len(cfg["upsample_rates"])
            if last and fused_post:
                s = f32(F.leaky_relu(f32(xs * (1.0 / nk)), 0.01))
                break
            s = stream(xs * (1.0 / nk), 0.01 if last else LRELU_SLOPE)
            if taps is not None:
                taps[f"mrf_act.{i}"] = s.float()
        x = f32(F.conv1d(s, (sd_f32 if fused_post else W)["conv_post.weight"], W["conv_post.bias"], padding=3))
        x = torch.tanh(x)
    return x.float()
