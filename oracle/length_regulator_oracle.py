"""CPU oracle for the FastSpeech2-style LengthRegulator (TEST INFRASTRUCTURE).

numpy restatement of scripts/model/modules.py:132-159 (LR / expand / forward) and of
``pad`` (scripts/utils/tools.py:669-687), plus ``get_mask_from_lengths``
(scripts/utils/tools.py:164-171) and the duration rounding rule
(scripts/model/modules.py:110-113).  Pinned by tests/golden/length_regulator_*.npz, which
tests/golden/make_golden.py produced by running the reference class itself.
"""
from __future__ import annotations

import numpy as np


def expand_counts(duration: np.ndarray) -> np.ndarray:
    """n_i = max(int(d_i), 0): python int() truncates toward zero (modules.py:152-153).

    NaN/inf would raise in the reference (int(nan)); not part of the contract.
    """
    d = np.asarray(duration)
    if d.dtype.kind == "f":
        d = np.trunc(d)
    return np.maximum(d.astype(np.int64), 0)


def length_regulate(x: np.ndarray, duration: np.ndarray, max_len=None):
    """x [B,T,H], duration [B,T] -> (out [B,Lmax,H], mel_len int64 [B]).

    modules.py:132-145: per row expand; tools.py:669-687: Lmax = max_len if truthy else
    max row length; rows longer than Lmax are cropped (negative F.pad), shorter ones
    zero-padded; mel_len reports the UNCROPPED length.
    """
    x = np.asarray(x)
    n = expand_counts(duration)
    B, T, H = x.shape
    mel_len = n.sum(axis=1).astype(np.int64)
    Lmax = int(max_len) if max_len else (int(mel_len.max()) if B else 0)
    out = np.zeros((B, Lmax, H), dtype=x.dtype)
    for b in range(B):
        idx = np.repeat(np.arange(T), n[b])[:Lmax]
        out[b, : idx.shape[0]] = x[b, idx]
    return out, mel_len


def get_mask_from_lengths(lengths: np.ndarray, max_len=None) -> np.ndarray:
    """tools.py:164-171: mask[b,p] = p >= lengths[b]; width max(lengths) unless given."""
    lengths = np.asarray(lengths)
    if max_len is None:
        max_len = int(lengths.max())
    ids = np.arange(max_len)[None, :]
    return ids >= lengths[:, None]


def round_durations(log_d: np.ndarray, d_control: float = 1.0) -> np.ndarray:
    """modules.py:110-113: clamp(round(exp(logd) - 1) * d_control, min=0); round half-to-even."""
    log_d = np.asarray(log_d, dtype=np.float32)
    r = np.round(np.exp(log_d) - np.float32(1.0)).astype(np.float32)   # np.round = half-to-even
    return np.maximum(r * np.float32(d_control), np.float32(0.0))
