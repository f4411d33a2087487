"""CPU restatement of PostNet.forward and mel_linear (TEST INFRASTRUCTURE -- only tests/, smoke() and bench.py's CPU
arm may import this; the product path is libvo2w).

Follows scripts/transformer/Layers.py:126-137 in eval mode: transpose, tanh(BatchNorm1d(Conv1d)) for all but the
last convolution, BatchNorm1d(Conv1d) for the last, transpose; F.dropout(..., training=False) is the identity.
BatchNorm1d in eval mode is y = (x - running_mean) / sqrt(running_var + eps) * weight + bias (eps = 1e-5), applied
here UNFOLDED, so the fold done by the drop-in is checked against it.  Pinned by tests/golden/postnet.npz, generated
from the imported reference class by tests/golden/make_golden.py.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

BN_EPS = 1e-5


def postnet_forward(sd, x: torch.Tensor, n_conv: int = 5) -> torch.Tensor:
    """sd: reference state dict (convolutions.{i}.0.conv.*, convolutions.{i}.1.*); x [B,L,n_mel] fp32."""
    with torch.no_grad():
        h = x.contiguous().transpose(1, 2)
        for i in range(n_conv):
            w, b = sd[f"convolutions.{i}.0.conv.weight"], sd[f"convolutions.{i}.0.conv.bias"]
            h = F.conv1d(h, w, b, stride=1, padding=(w.shape[2] - 1) // 2, dilation=1)
            h = F.batch_norm(h, sd[f"convolutions.{i}.1.running_mean"], sd[f"convolutions.{i}.1.running_var"],
                             sd[f"convolutions.{i}.1.weight"], sd[f"convolutions.{i}.1.bias"], False, 0.1, BN_EPS)
            if i < n_conv - 1:
                h = torch.tanh(h)
        return h.contiguous().transpose(1, 2)


def vtts_tail(sd, lin_w, lin_b, dec_out: torch.Tensor):
    """scripts/model/vtts.py:106-107: output = mel_linear(dec_out); postnet_output = postnet(output) + output."""
    with torch.no_grad():
        output = F.linear(dec_out, lin_w, lin_b)
        return output, postnet_forward(sd, output) + output
