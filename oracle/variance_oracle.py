"""CPU restatement of VariancePredictor / VarianceAdaptor (TEST INFRASTRUCTURE -- only tests/, smoke() and bench.py's
CPU arm may import this; the product path is libvo2w).

Follows scripts/model/modules.py in eval mode (dropout is the identity):
  VariancePredictor.forward      modules.py:203-213  (conv_layer Sequential: modules.py:172-199; Conv: modules.py:256-262)
  get_energy_embedding           modules.py:53-65    (kurtosis: 67-78)
  VarianceAdaptor.forward        modules.py:80-124
with functional ops on the reference's own state-dict keys; LengthRegulator and get_mask_from_lengths are the numpy
restatements of length_regulator_oracle.py.  Pinned by tests/golden/variance.npz, generated from the imported reference
classes by tests/golden/make_golden.py.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

from . import length_regulator_oracle as lro

LN_EPS = 1e-5


def variance_predictor(sd, pre, x, mask):
    with torch.no_grad():
        h = x
        for i in (1, 2):
            w, b = sd[f"{pre}conv_layer.conv1d_{i}.conv.weight"], sd[f"{pre}conv_layer.conv1d_{i}.conv.bias"]
            pad = (w.shape[2] - 1) // 2 if i == 1 else 1
            h = F.conv1d(h.contiguous().transpose(1, 2), w, b, padding=pad).contiguous().transpose(1, 2)
            h = F.layer_norm(F.relu(h), (h.shape[2],), sd[f"{pre}conv_layer.layer_norm_{i}.weight"],
                             sd[f"{pre}conv_layer.layer_norm_{i}.bias"], LN_EPS)
        out = F.linear(h, sd[pre + "linear_layer.weight"], sd[pre + "linear_layer.bias"]).squeeze(-1)
        if mask is not None:
            out = out.masked_fill(mask, 0.0)
        return out


def variance_embedding(sd, name, prediction, target, mean, std, control):
    """name: 'energy' or 'kurt'.  Returns (prediction, embedding)."""
    bins, table = sd[f"{name}_bins"], sd[f"{name}_embedding.weight"]
    if target is not None:
        return prediction, F.embedding(torch.bucketize(target, bins), table)
    prediction = prediction * std + mean
    prediction = prediction * control
    prediction = (prediction - mean) / std
    return prediction, F.embedding(torch.bucketize(prediction, bins), table)


def variance_adaptor_forward(sd, stats, x, src_mask, max_len=None, energy_target=None, kurtosis_target=None,
                             duration_target=None, e_control=1.0, d_control=1.0, is_energy=True, is_kurtosis=False):
    """stats = {"energy": [min, max, mean, std], "kurtosis": [...]} (stats.json).  Returns the reference's 7-tuple."""
    with torch.no_grad():
        log_d = variance_predictor(sd, "duration_predictor.", x, src_mask)
        e_pred = k_pred = None
        if is_energy:
            e_pred = variance_predictor(sd, "energy_predictor.", x, src_mask)
            e_pred, emb = variance_embedding(sd, "energy", e_pred, energy_target, stats["energy"][2], stats["energy"][3], e_control)
            x = x + emb
        if is_kurtosis:
            k_pred = variance_predictor(sd, "kurtosis_predictor.", x, src_mask)
            k_pred, emb = variance_embedding(sd, "kurt", k_pred, kurtosis_target, stats["kurtosis"][2], stats["kurtosis"][3], 1.0)
            x = x + emb
        mel_mask = None
        if duration_target is not None:
            d_round = duration_target
        else:
            d_round = torch.clamp(torch.round(torch.exp(log_d) - 1) * d_control, min=0)
        out, mel_len = lro.length_regulate(x.numpy(), d_round.numpy(), max_len)
        if duration_target is None:
            mel_mask = torch.from_numpy(lro.get_mask_from_lengths(mel_len))
        return torch.from_numpy(out), e_pred, k_pred, log_d, d_round, torch.from_numpy(mel_len), mel_mask
