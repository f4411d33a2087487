"""Drop-in for the reference's LengthRegulator (scripts/model/modules.py:126-159).

The reference expands token by token in Python with one ``.item()`` device->host sync per
token (B*T_src syncs) plus ``pad`` (scripts/utils/tools.py:669-687).  Here one CUDA launch
(prefix sum + vectorised gather, csrc/lr.cu) does the expansion; at most ONE scalar is read
back, and only when the reference semantics need it (``max_len`` falsy => output width is the
longest row).  Bit-exact with the reference, including int() truncation of float
durations, negatives / -0.0 -> 0, cropping when ``max_len`` is shorter than a row while
``mel_len`` still reports the uncropped length, and the all-zero batch ([B,0,H]).
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _lib

_DUR = {torch.int32: _lib.DUR_I32, torch.int64: _lib.DUR_I64, torch.float32: _lib.DUR_F32,
        torch.float64: _lib.DUR_F64}


def _prep_duration(duration: torch.Tensor, device) -> torch.Tensor:
    d = duration.to(device)
    if d.dtype not in _DUR:
        d = d.to(torch.float32 if d.is_floating_point() else torch.int64)   # exact for fp16/bf16/int8/int16/bool
    return d.contiguous()


class LengthRegulator(nn.Module):
    """Length Regulator (scripts/model/modules.py:126-159)."""

    def __init__(self):
        super().__init__()

    def LR(self, x, duration, max_len, return_mask: bool = False):
        if not isinstance(x, torch.Tensor) or x.dim() != 3:
            raise RuntimeError("LengthRegulator expects x [B,T_src,H]")
        if not x.is_cuda:
            raise RuntimeError("LengthRegulator: x must be a CUDA tensor (this path has no CPU fallback)")
        B, T, H = x.shape
        if B == 0:
            raise ValueError("max() arg is an empty sequence")     # what tools.pad raises on an empty batch
        if tuple(duration.shape) != (B, T):
            raise RuntimeError(f"duration shape {tuple(duration.shape)} does not match x {tuple(x.shape)}")
        lib = _lib.lib()
        x = x.contiguous()
        with torch.cuda.device(x.device):
            d = _prep_duration(duration, x.device)
            stream = torch.cuda.current_stream(x.device).cuda_stream
            mel_len = torch.empty(B, dtype=torch.int64, device=x.device)
            if max_len:
                Lmax = int(max_len)
            else:
                mx = torch.empty(1, dtype=torch.int64, device=x.device)
                _lib.check(lib.vo2w_lr_lengths(d.data_ptr(), _DUR[d.dtype], B, T, mel_len.data_ptr(), mx.data_ptr(),
                                               stream), "vo2w_lr_lengths")
                Lmax = int(mx.item())         # the single device->host read (reference: B*T_src of them)
            out = torch.empty(B, Lmax, H, dtype=x.dtype, device=x.device)
            mask = torch.empty(B, Lmax, dtype=torch.bool, device=x.device) if return_mask else None
            _lib.check(lib.vo2w_length_regulate(x.data_ptr(), d.data_ptr(), _DUR[d.dtype], B, T,
                                                H * x.element_size(), out.data_ptr(), Lmax, mel_len.data_ptr(),
                                                mask.data_ptr() if mask is not None else None, stream),
                       "vo2w_length_regulate")
        if return_mask:
            return out, mel_len, mask
        return out, mel_len

    def forward(self, x, duration, max_len):
        output, mel_len = self.LR(x, duration, max_len)
        return output, mel_len


def get_mask_from_lengths(lengths: torch.Tensor, max_len=None) -> torch.Tensor:
    """scripts/utils/tools.py:164-171: mask[b,p] = p >= lengths[b]; width max(lengths) unless given."""
    if not lengths.is_cuda:
        raise RuntimeError("get_mask_from_lengths: lengths must be a CUDA tensor")
    B = lengths.shape[0]
    lengths = lengths.to(torch.int64).contiguous()
    if max_len is None:
        max_len = int(torch.max(lengths).item())
    mask = torch.empty(B, int(max_len), dtype=torch.bool, device=lengths.device)
    with torch.cuda.device(lengths.device):
        stream = torch.cuda.current_stream(lengths.device).cuda_stream
        _lib.check(_lib.lib().vo2w_mask_from_lengths(lengths.data_ptr(), B, int(max_len), mask.data_ptr(), stream),
                   "vo2w_mask_from_lengths")
    return mask


def round_durations(log_duration_prediction: torch.Tensor, d_control: float = 1.0) -> torch.Tensor:
    """The duration rounding rule of VarianceAdaptor (scripts/model/modules.py:110-113); stock torch."""
    return torch.clamp(torch.round(torch.exp(log_duration_prediction) - 1) * d_control, min=0)
