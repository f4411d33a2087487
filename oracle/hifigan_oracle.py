"""CPU oracle for the HiFi-GAN V1 Generator (TEST INFRASTRUCTURE, not product code).

Restates ``Generator.forward`` of the reference (scripts/hifigan/models.py:149-165) and
``ResBlock.forward`` (scripts/hifigan/models.py:96-103) as a pure function over a folded
state-dict, using the same library calls the reference dispatches (the reference's
arithmetic lives in PyTorch: F.conv1d / F.conv_transpose1d / F.leaky_relu / torch.tanh),
so the fp32 result is the reference's result.  ``fold_weight_norm`` restates
``remove_weight_norm`` (scripts/hifigan/models.py:105-109,167-174).

Pinned by tests/golden/generator_*.npz (made by tests/golden/make_golden.py from the
imported reference); the reference has no golden vectors of its own.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch
import torch.nn.functional as F

LRELU_SLOPE = 0.1          # scripts/hifigan/models.py:7
FINAL_LRELU_SLOPE = 0.01   # F.leaky_relu default, scripts/hifigan/models.py:161

# scripts/hifigan/config.json:11-15 (HiFi-GAN V1)
V1_CONFIG = {
    "upsample_rates": [8, 8, 2, 2],
    "upsample_kernel_sizes": [16, 16, 4, 4],
    "upsample_initial_channel": 512,
    "resblock_kernel_sizes": [3, 7, 11],
    "resblock_dilation_sizes": [[1, 3, 5], [1, 3, 5], [1, 3, 5]],
    "sampling_rate": 22050,
    "hop_size": 256,
}

FLOP_PER_FRAME = 614_105_088  # SURVEY.md section 8(d); checked in tests against count_flops()


def get_padding(kernel_size: int, dilation: int = 1) -> int:
    """scripts/hifigan/models.py:16-17."""
    return int((kernel_size * dilation - dilation) / 2)


def layer_names(cfg=V1_CONFIG):
    """All 78 conv layer prefixes in state-dict order (scripts/hifigan/models.py:118-145)."""
    names = ["conv_pre"]
    names += [f"ups.{i}" for i in range(len(cfg["upsample_rates"]))]
    nk = len(cfg["resblock_kernel_sizes"])
    for i in range(len(cfg["upsample_rates"])):
        for j in range(nk):
            for grp in ("convs1", "convs2"):
                for m in range(3):
                    names.append(f"resblocks.{i * nk + j}.{grp}.{m}")
    names.append("conv_post")
    return names


def layer_shapes(cfg=V1_CONFIG):
    """name -> (weight shape, is_transposed).  Mirrors the constructor at models.py:113-145."""
    c0 = cfg["upsample_initial_channel"]
    shapes = {"conv_pre": ((c0, 80, 7), False)}
    nk = len(cfg["resblock_kernel_sizes"])
    ch = c0
    for i, (u, k) in enumerate(zip(cfg["upsample_rates"], cfg["upsample_kernel_sizes"])):
        shapes[f"ups.{i}"] = ((c0 // 2 ** i, c0 // 2 ** (i + 1), k), True)
        ch = c0 // 2 ** (i + 1)
        for j, rk in enumerate(cfg["resblock_kernel_sizes"]):
            for grp in ("convs1", "convs2"):
                for m in range(3):
                    shapes[f"resblocks.{i * nk + j}.{grp}.{m}"] = ((ch, ch, rk), False)
    shapes["conv_post"] = ((1, ch, 7), False)
    return shapes


def fold_weight_norm(sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """234-key {bias, weight_g, weight_v} -> 156-key {weight, bias}.

    ``remove_weight_norm`` (models.py:105-109,167-174) replaces g,v by
    ``torch._weight_norm(v, g, dim=0)``; dim 0 is Cout for Conv1d and **Cin** for
    ConvTranspose1d.  Already-folded dicts pass through.
    """
    out: Dict[str, torch.Tensor] = {}
    for k, v in sd.items():
        if k.endswith(".weight_g"):
            base = k[: -len(".weight_g")]
            out[base + ".weight"] = torch._weight_norm(sd[base + ".weight_v"], v, 0)
        elif k.endswith(".weight_v"):
            continue
        else:
            out[k] = v
    return out


def resblock_forward(sd, prefix: str, x: torch.Tensor, k: int, dilations) -> torch.Tensor:
    """scripts/hifigan/models.py:96-103."""
    for m, d in enumerate(dilations):
        xt = F.leaky_relu(x, LRELU_SLOPE)
        xt = F.conv1d(xt, sd[f"{prefix}.convs1.{m}.weight"], sd[f"{prefix}.convs1.{m}.bias"],
                      stride=1, padding=get_padding(k, d), dilation=d)
        xt = F.leaky_relu(xt, LRELU_SLOPE)
        xt = F.conv1d(xt, sd[f"{prefix}.convs2.{m}.weight"], sd[f"{prefix}.convs2.{m}.bias"],
                      stride=1, padding=get_padding(k, 1), dilation=1)
        x = xt + x
    return x


def generator_forward(sd: Dict[str, torch.Tensor], mel: torch.Tensor, cfg=V1_CONFIG,
                      taps: Optional[dict] = None) -> torch.Tensor:
    """mel [B,80,T] -> wav [B,1,256T].  scripts/hifigan/models.py:149-165.

    ``sd`` is a folded (or g/v) state-dict; dtype of ``mel`` selects fp32 or the fp64
    arbiter (weights are cast to it).  ``taps``, if given, is filled with intermediate
    activations (per-layer parity / debugging aid).
    """
    sd = fold_weight_norm(sd)
    sd = {k: v.to(mel.dtype) for k, v in sd.items()}
    nk = len(cfg["resblock_kernel_sizes"])
    with torch.no_grad():
        x = F.conv1d(mel, sd["conv_pre.weight"], sd["conv_pre.bias"], stride=1, padding=3)
        if taps is not None:
            taps["conv_pre"] = x
        for i, (u, k) in enumerate(zip(cfg["upsample_rates"], cfg["upsample_kernel_sizes"])):
            x = F.leaky_relu(x, LRELU_SLOPE)
            x = F.conv_transpose1d(x, sd[f"ups.{i}.weight"], sd[f"ups.{i}.bias"],
                                   stride=u, padding=(k - u) // 2)
            if taps is not None:
                taps[f"ups.{i}"] = x
            xs = None
            for j, (rk, rd) in enumerate(zip(cfg["resblock_kernel_sizes"],
                                             cfg["resblock_dilation_sizes"])):
                r = resblock_forward(sd, f"resblocks.{i * nk + j}", x, rk, rd)
                if taps is not None:
                    taps[f"resblocks.{i * nk + j}"] = r
                if xs is None:
                    xs = r
                else:
                    xs += r
            x = xs / nk
            if taps is not None:
                taps[f"mrf.{i}"] = x
        x = F.leaky_relu(x)            # slope 0.01 (models.py:161)
        x = F.conv1d(x, sd["conv_post.weight"], sd["conv_post.bias"], stride=1, padding=3)
        x = torch.tanh(x)
    return x


def count_flops(T: int = 1, B: int = 1, cfg=V1_CONFIG) -> int:
    """Analytic FLOPs (2 x MAC) of generator_forward; 614 105 088 per frame for V1."""
    macs = 0
    L = T
    for name, (shape, transposed) in layer_shapes(cfg).items():
        if name.startswith("ups."):
            i = int(name.split(".")[1])
            cin, cout, k = shape
            macs += L * cin * cout * k       # every input sample hits all k taps
            L *= cfg["upsample_rates"][i]
        else:
            cout, cin, k = shape
            macs += L * cin * cout * k
    return 2 * macs * B
