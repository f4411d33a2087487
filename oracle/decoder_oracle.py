"""CPU restatement of the mel Decoder's FFT-block stack (TEST INFRASTRUCTURE -- only tests/, smoke() and bench.py's
CPU arm may import this; the product path is libvo2w).

Follows, in eval mode (every dropout is the identity):
  Decoder.forward                  scripts/transformer/Models.py:168-197
  FFTBlock.forward                 scripts/transformer/Layers.py:21-30
  MultiHeadAttention.forward       scripts/transformer/SubLayers.py:31-62
  ScaledDotProductAttention        scripts/transformer/Modules.py:15-25
  PositionwiseFeedForward.forward  scripts/transformer/SubLayers.py:95-104
  get_sinusoid_encoding_table      scripts/transformer/Models.py:13-33
written with functional ops on the reference's own state-dict keys.  Pinned by tests/golden/decoder.npz, generated
from the imported reference class by tests/golden/make_golden.py.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

LN_EPS = 1e-5


def sinusoid_table(n_position: int, d_hid: int) -> torch.Tensor:
    table = np.array([[pos / np.power(10000, 2 * (j // 2) / d_hid) for j in range(d_hid)] for pos in range(n_position)])
    table[:, 0::2] = np.sin(table[:, 0::2])
    table[:, 1::2] = np.cos(table[:, 1::2])
    return torch.FloatTensor(table)


def multi_head_attention(sd, pre, x, key_mask, n_head):
    B, L, D = x.shape
    dk = D // n_head
    def heads(name):
        y = F.linear(x, sd[pre + name + ".weight"], sd[pre + name + ".bias"]).view(B, L, n_head, dk)
        return y.permute(2, 0, 1, 3).contiguous().view(-1, L, dk)                  # (n*b) x L x dk
    q, k, v = heads("w_qs"), heads("w_ks"), heads("w_vs")
    attn = torch.bmm(q, k.transpose(1, 2)) / np.power(dk, 0.5)
    m = key_mask.unsqueeze(1).expand(-1, L, -1).repeat(n_head, 1, 1)
    attn = torch.softmax(attn.masked_fill(m, -np.inf), dim=2)
    out = torch.bmm(attn, v).view(n_head, B, L, dk).permute(1, 2, 0, 3).contiguous().view(B, L, -1)
    out = F.linear(out, sd[pre + "fc.weight"], sd[pre + "fc.bias"])
    return F.layer_norm(out + x, (D,), sd[pre + "layer_norm.weight"], sd[pre + "layer_norm.bias"], LN_EPS)


def positionwise_ffn(sd, pre, x):
    w1, w2 = sd[pre + "w_1.weight"], sd[pre + "w_2.weight"]
    h = F.conv1d(x.transpose(1, 2), w1, sd[pre + "w_1.bias"], padding=(w1.shape[2] - 1) // 2)
    h = F.conv1d(F.relu(h), w2, sd[pre + "w_2.bias"], padding=(w2.shape[2] - 1) // 2).transpose(1, 2)
    return F.layer_norm(h + x, (x.shape[2],), sd[pre + "layer_norm.weight"], sd[pre + "layer_norm.bias"], LN_EPS)


def decoder_forward(sd, enc_seq: torch.Tensor, mask: torch.Tensor, n_head: int = 2, max_seq_len: int | None = None):
    """sd: reference state dict (position_enc, layer_stack.{i}.*); enc_seq [B,L,D] fp32; mask [B,L] bool (True = padding).
    Returns (dec_output [B,L,D], mask) like the reference in eval mode."""
    with torch.no_grad():
        B, L, D = enc_seq.shape
        table = sd["position_enc"][0]
        if max_seq_len is None:
            max_seq_len = table.shape[0] - 1
        pos = sinusoid_table(L, D) if L > max_seq_len else table[:L]
        x = enc_seq + pos.unsqueeze(0)
        n_layers = 1 + max(int(k.split(".")[1]) for k in sd if k.startswith("layer_stack."))
        for i in range(n_layers):
            x = multi_head_attention(sd, f"layer_stack.{i}.slf_attn.", x, mask, n_head)
            x = x.masked_fill(mask.unsqueeze(-1), 0)
            x = positionwise_ffn(sd, f"layer_stack.{i}.pos_ffn.", x)
            x = x.masked_fill(mask.unsqueeze(-1), 0)
        return x, mask
