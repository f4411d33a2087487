"""Deterministic synthetic weights / inputs for benchmarks, smoke and tests.

No checkpoint ships with the reference (generator_universal.pth.tar is a missing blob,
SURVEY.md facts table), so every number in this repo is on random-init weights.  The
reference's random init is nn.Conv1d's default, U(-1/sqrt(fan_in), 1/sqrt(fan_in)):
``init_weights`` (scripts/hifigan/models.py:10-13) writes into the recomputed ``.weight``
of a weight_norm module and is lost.  This module reproduces those magnitudes without
touching the reference, so the same weights exist on the GPU box.
"""
from __future__ import annotations

import math
from typing import Dict

import torch

V1_CONFIG = {  # scripts/hifigan/config.json:11-15,21,24
    "upsample_rates": [8, 8, 2, 2],
    "upsample_kernel_sizes": [16, 16, 4, 4],
    "upsample_initial_channel": 512,
    "resblock_kernel_sizes": [3, 7, 11],
    "resblock_dilation_sizes": [[1, 3, 5], [1, 3, 5], [1, 3, 5]],
    "sampling_rate": 22050,
    "hop_size": 256,
}

FLOP_PER_FRAME = 614_105_088
# the 72 ResBlock convolutions alone (2 x (66 060 288 + 132 120 576 + 66 060 288 + 33 030 144) MAC)
PAIR_FLOP_PER_FRAME = 594_542_592


def layer_shapes(cfg=V1_CONFIG):
    """name -> (weight shape, is ConvTranspose1d), in state-dict order (models.py:113-145)."""
    c0 = cfg["upsample_initial_channel"]
    shapes = {"conv_pre": ((c0, 80, 7), False)}
    nk = len(cfg["resblock_kernel_sizes"])
    ch = c0
    for i, k in enumerate(cfg["upsample_kernel_sizes"]):
        shapes[f"ups.{i}"] = ((c0 >> i, c0 >> (i + 1), k), True)
    for i in range(len(cfg["upsample_rates"])):
        ch = c0 >> (i + 1)
        for j, rk in enumerate(cfg["resblock_kernel_sizes"]):
            for grp in ("convs1", "convs2"):
                for m in range(3):
                    shapes[f"resblocks.{i * nk + j}.{grp}.{m}"] = ((ch, ch, rk), False)
    shapes["conv_post"] = ((1, ch, 7), False)
    return shapes


def seeded_state_dict(seed: int, cfg=V1_CONFIG, weight_norm: bool = False) -> Dict[str, torch.Tensor]:
    """234-key {bias, weight_g, weight_v} (weight_norm=True, g = ||v||) or 156-key folded."""
    gen = torch.Generator().manual_seed(seed)
    sd: Dict[str, torch.Tensor] = {}
    for name, (shape, transposed) in layer_shapes(cfg).items():
        bound = 1.0 / math.sqrt(shape[1] * shape[2])   # torch fan_in = size(1) * k
        w = (torch.rand(shape, generator=gen) * 2 - 1) * bound
        nb = shape[1] if transposed else shape[0]
        b = (torch.rand(nb, generator=gen) * 2 - 1) * bound
        sd[f"{name}.bias"] = b
        if weight_norm:
            sd[f"{name}.weight_g"] = w.reshape(shape[0], -1).norm(dim=1).reshape(shape[0], 1, 1)
            sd[f"{name}.weight_v"] = w
        else:
            sd[f"{name}.weight"] = w
    return sd


def synthetic_mel(B: int, T: int, seed: int = 0, kind: str = "logmel") -> torch.Tensor:
    """[B,80,T] fp32.  'logmel' = rand*13.5-11.5 (floor ln 1e-5)."""
    gen = torch.Generator().manual_seed(seed)
    if kind == "logmel":
        return torch.rand(B, 80, T, generator=gen) * 13.5 - 11.5
    return torch.randn(B, 80, T, generator=gen)


def synthetic_durations(B: int, T_src: int, seed: int = 0, hi: int = 25) -> torch.Tensor:
    """int64 [B,T_src] ~ U{0..hi-1} (config 3)."""
    gen = torch.Generator().manual_seed(seed)
    return torch.randint(0, hi, (B, T_src), generator=gen)


def seeded_postnet_state_dict(seed: int = 4321, n_mel: int = 80, dim: int = 512, k: int = 5, n_conv: int = 5):
    """State dict with the reference PostNet's keys (scripts/transformer/Layers.py:67-124): seeded convolution
    weights and NON-trivial BatchNorm statistics, so that the eval-mode fold is actually exercised."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    chans = [n_mel] + [dim] * (n_conv - 1) + [n_mel]
    for i in range(n_conv):
        cin, cout = chans[i], chans[i + 1]
        sd[f"convolutions.{i}.0.conv.weight"] = torch.randn(cout, cin, k, generator=g) * (1.5 / (cin * k) ** 0.5)
        sd[f"convolutions.{i}.0.conv.bias"] = torch.randn(cout, generator=g) * 0.1
        sd[f"convolutions.{i}.1.weight"] = torch.rand(cout, generator=g) + 0.5
        sd[f"convolutions.{i}.1.bias"] = torch.randn(cout, generator=g) * 0.1
        sd[f"convolutions.{i}.1.running_mean"] = torch.randn(cout, generator=g) * 0.2
        sd[f"convolutions.{i}.1.running_var"] = torch.rand(cout, generator=g) + 0.5
        sd[f"convolutions.{i}.1.num_batches_tracked"] = torch.tensor(100)
    return sd


def sinusoid_table(n_position: int, d_hid: int) -> torch.Tensor:
    """The reference's position table (scripts/transformer/Models.py:13-33): float64 angles
    pos / 10000^(2 (j // 2) / d_hid), sin on even columns, cos on odd ones, cast to float32."""
    import numpy as np
    pos = np.arange(n_position, dtype=np.float64)[:, None]
    j = np.arange(d_hid)[None, :]
    table = pos / np.power(10000, 2 * (j // 2) / d_hid)
    table[:, 0::2] = np.sin(table[:, 0::2])
    table[:, 1::2] = np.cos(table[:, 1::2])
    return torch.FloatTensor(table)


def seeded_decoder_state_dict(seed: int = 2468, d_model: int = 256, n_head: int = 2, d_inner: int = 1024, kernels=(9, 1),
                              n_layers: int = 6, max_seq_len: int = 1000):
    """State dict with the reference Decoder's keys (scripts/transformer/Models.py:137-166, SubLayers.py): seeded
    Linear / Conv1d weights and NON-trivial LayerNorm affine parameters."""
    g = torch.Generator().manual_seed(seed)
    sd = {"position_enc": sinusoid_table(max_seq_len + 1, d_model).unsqueeze(0)}
    for i in range(n_layers):
        a, f = f"layer_stack.{i}.slf_attn.", f"layer_stack.{i}.pos_ffn."
        for name in ("w_qs", "w_ks", "w_vs", "fc"):
            sd[a + name + ".weight"] = torch.randn(d_model, d_model, generator=g) * (1.2 / d_model ** 0.5)
            sd[a + name + ".bias"] = torch.randn(d_model, generator=g) * 0.1
        sd[a + "layer_norm.weight"] = torch.rand(d_model, generator=g) + 0.5
        sd[a + "layer_norm.bias"] = torch.randn(d_model, generator=g) * 0.1
        sd[f + "w_1.weight"] = torch.randn(d_inner, d_model, kernels[0], generator=g) * (1.4 / (d_model * kernels[0]) ** 0.5)
        sd[f + "w_1.bias"] = torch.randn(d_inner, generator=g) * 0.1
        sd[f + "w_2.weight"] = torch.randn(d_model, d_inner, kernels[1], generator=g) * (1.4 / (d_inner * kernels[1]) ** 0.5)
        sd[f + "w_2.bias"] = torch.randn(d_model, generator=g) * 0.1
        sd[f + "layer_norm.weight"] = torch.rand(d_model, generator=g) + 0.5
        sd[f + "layer_norm.bias"] = torch.randn(d_model, generator=g) * 0.1
    return sd


def decoder_config(d_model: int = 256, n_head: int = 2, d_inner: int = 1024, kernels=(9, 1), n_layers: int = 6,
                   max_seq_len: int = 1000):
    """The slice of config/*/model.yaml Decoder.__init__ reads (Models.py:140-154)."""
    return {"max_seq_len": max_seq_len,
            "transformer": {"decoder_hidden": d_model, "decoder_layer": n_layers, "decoder_head": n_head,
                            "conv_filter_size": d_inner, "conv_kernel_size": list(kernels), "decoder_dropout": 0.2}}


VARIANCE_STATS = {"energy": [-1.431, 8.012, 37.32, 26.08], "kurtosis": [-1.207, 9.034, 3.127, 2.215]}   # a stats.json


def variance_model_config(d_hidden: int = 256, filter_size: int = 256, kernel: int = 3, n_bins: int = 256,
                          is_energy: bool = True, is_kurtosis: bool = False):
    """The slice of config/*/model.yaml VarianceAdaptor / VariancePredictor read (modules.py:17-50, 164-170)."""
    return {"transformer": {"encoder_hidden": d_hidden},
            "variance_predictor": {"filter_size": filter_size, "kernel_size": kernel, "dropout": 0.5},
            "variance_embedding": {"is_kurtosis_condition": is_kurtosis, "is_energy_condition": is_energy,
                                   "kurtosis_quantization": "linear", "energy_quantization": "linear", "n_bins": n_bins}}


def seeded_variance_adaptor_state_dict(seed: int = 1357, d_hidden: int = 256, filter_size: int = 256, kernel: int = 3,
                                       n_bins: int = 256, is_energy: bool = True, is_kurtosis: bool = False,
                                       stats=VARIANCE_STATS):
    """State dict with the reference VarianceAdaptor's keys (scripts/model/modules.py:17-50): seeded predictors whose
    outputs spread over the bins / over durations 0..~30, linear bin edges from `stats`, seeded embedding tables."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    preds = ["duration_predictor"] + (["kurtosis_predictor"] if is_kurtosis else []) + (["energy_predictor"] if is_energy else [])
    for name in preds:
        p = name + "."
        sd[p + "conv_layer.conv1d_1.conv.weight"] = torch.randn(filter_size, d_hidden, kernel, generator=g) * (1.4 / (d_hidden * kernel) ** 0.5)
        sd[p + "conv_layer.conv1d_1.conv.bias"] = torch.randn(filter_size, generator=g) * 0.1
        sd[p + "conv_layer.layer_norm_1.weight"] = torch.rand(filter_size, generator=g) + 0.5
        sd[p + "conv_layer.layer_norm_1.bias"] = torch.randn(filter_size, generator=g) * 0.1
        sd[p + "conv_layer.conv1d_2.conv.weight"] = torch.randn(filter_size, filter_size, kernel, generator=g) * (1.4 / (filter_size * kernel) ** 0.5)
        sd[p + "conv_layer.conv1d_2.conv.bias"] = torch.randn(filter_size, generator=g) * 0.1
        sd[p + "conv_layer.layer_norm_2.weight"] = torch.rand(filter_size, generator=g) + 0.5
        sd[p + "conv_layer.layer_norm_2.bias"] = torch.randn(filter_size, generator=g) * 0.1
        sd[p + "linear_layer.weight"] = torch.randn(1, filter_size, generator=g) * 0.06
        sd[p + "linear_layer.bias"] = torch.tensor([1.2 if name == "duration_predictor" else 0.3])
    sd["kurt_bins"] = torch.linspace(stats["kurtosis"][0], stats["kurtosis"][1], n_bins - 1)
    sd["kurt_embedding.weight"] = torch.randn(n_bins, d_hidden, generator=g)
    sd["energy_bins"] = torch.linspace(stats["energy"][0], stats["energy"][1], n_bins - 1)
    sd["energy_embedding.weight"] = torch.randn(n_bins, d_hidden, generator=g)
    return sd
