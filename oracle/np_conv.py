"""numpy restatement of the two convolution primitives (TEST INFRASTRUCTURE).

Independent of torch: pins the meaning of the ATen calls the reference makes
(nn.Conv1d at scripts/hifigan/models.py:26-92,118-120,145 and nn.ConvTranspose1d at
models.py:123-135) and states the polyphase identity the CUDA upsampler relies on
(SURVEY.md section 8 a3).  Small cases only.
"""
import numpy as np


def conv1d_np(x, w, b, padding, dilation=1):
    """x [B,Cin,L], w [Cout,Cin,k] -> [B,Cout,L + 2p - d(k-1)] (cross-correlation)."""
    B, Cin, L = x.shape
    Cout, _, k = w.shape
    xp = np.pad(x, ((0, 0), (0, 0), (padding, padding)))
    Lo = L + 2 * padding - dilation * (k - 1)
    y = np.zeros((B, Cout, Lo), dtype=x.dtype)
    for j in range(k):
        y += np.einsum("oc,bcl->bol", w[:, :, j], xp[:, :, j * dilation: j * dilation + Lo])
    return y + b[None, :, None]


def conv_transpose1d_np(x, w, b, stride, padding):
    """x [B,Cin,L], w [Cin,Cout,k] -> [B,Cout,(L-1)s - 2p + k]; out[s*i + j - p] += x[i] w[j]."""
    B, Cin, L = x.shape
    _, Cout, k = w.shape
    full = np.zeros((B, Cout, (L - 1) * stride + k), dtype=x.dtype)
    for i in range(L):
        for j in range(k):
            full[:, :, i * stride + j] += np.einsum("bc,co->bo", x[:, :, i], w[:, :, j])
    Lo = (L - 1) * stride - 2 * padding + k
    return full[:, :, padding: padding + Lo] + b[None, :, None]


def conv_transpose1d_polyphase_np(x, w, b, stride):
    """Same result for k = 2s, p = s/2 written as s phases of 2 taps each.

    out[s*q + r] = x[q + e] . w[:, :, j0] + x[q + e - 1] . w[:, :, j0 + s]
    with j0 = (r + p) mod s and e = (r + p) div s  (e = 0 for r < s/2, else 1).
    """
    B, Cin, L = x.shape
    _, Cout, k = w.shape
    s = stride
    assert k == 2 * s and s % 2 == 0
    p = s // 2
    xp = np.pad(x, ((0, 0), (0, 0), (1, 1)))          # x[-1] = x[L] = 0
    y = np.zeros((B, Cout, s * L), dtype=x.dtype)
    for r in range(s):
        j0, e = (r + p) % s, (r + p) // s
        hi = xp[:, :, 1 + e: 1 + e + L]                # x[q + e]
        lo = xp[:, :, e: e + L]                        # x[q + e - 1]
        y[:, :, r::s] = (np.einsum("bcl,co->bol", hi, w[:, :, j0]) +
                         np.einsum("bcl,co->bol", lo, w[:, :, j0 + s]))
    return y + b[None, :, None]
