"""Drop-in for the acoustic model's ``VarianceAdaptor`` / ``VariancePredictor`` (scripts/model/modules.py:15-213):
the steps in front of LengthRegulator.

Same module tree and state-dict keys as the reference (``duration_predictor.conv_layer.conv1d_1.conv.weight``,
``...layer_norm_1.weight``, ``...linear_layer.weight``, ``energy_bins``, ``energy_embedding.weight``, ``kurt_bins``,
``kurt_embedding.weight`` ...), so a vTTS checkpoint's ``variance_adaptor.*`` entries load unchanged.  ``forward``
never runs a PyTorch layer: predictors, bucketize + embedding add, duration rounding, LengthRegulator and the mel mask
are libvo2w launches (csrc/variance.cu, csrc/lr.cu); the only device->host read is the one LengthRegulator needs for
its output width.  Inference only, CUDA only.
"""
from __future__ import annotations

import ctypes as C
import json
import os
from collections import OrderedDict

import numpy as np
import torch
import torch.nn as nn

from . import _lib
from .length_regulator import LengthRegulator, get_mask_from_lengths


class Conv(nn.Module):
    """scripts/model/modules.py:216-262: a Conv1d holder (``.conv``) applied on the time-major view."""

    def __init__(self, in_channels, out_channels, kernel_size=1, stride=1, padding=0, dilation=1, bias=True, w_init="linear"):
        super().__init__()
        self.conv = nn.Conv1d(in_channels, out_channels, kernel_size=kernel_size, stride=stride, padding=padding,
                              dilation=dilation, bias=bias)

    def forward(self, x):
        raise RuntimeError("Conv is a parameter holder on the B200 path; call VariancePredictor.forward")


class VariancePredictor(nn.Module):
    """Duration / energy / kurtosis predictor (modules.py:161-213)."""

    def __init__(self, model_config):
        super().__init__()
        self.input_size = model_config["transformer"]["encoder_hidden"]
        self.filter_size = model_config["variance_predictor"]["filter_size"]
        self.kernel = model_config["variance_predictor"]["kernel_size"]
        self.conv_output_size = model_config["variance_predictor"]["filter_size"]
        self.dropout = model_config["variance_predictor"]["dropout"]
        self.conv_layer = nn.Sequential(OrderedDict([
            ("conv1d_1", Conv(self.input_size, self.filter_size, kernel_size=self.kernel, padding=(self.kernel - 1) // 2)),
            ("relu_1", nn.ReLU()),
            ("layer_norm_1", nn.LayerNorm(self.filter_size)),
            ("dropout_1", nn.Dropout(self.dropout)),
            ("conv1d_2", Conv(self.filter_size, self.filter_size, kernel_size=self.kernel, padding=1)),
            ("relu_2", nn.ReLU()),
            ("layer_norm_2", nn.LayerNorm(self.filter_size)),
            ("dropout_2", nn.Dropout(self.dropout)),
        ]))
        self.linear_layer = nn.Linear(self.conv_output_size, 1)
        self._handle = None
        self._handle_dev = None
        self._ws = None

    def forward(self, encoder_output: torch.Tensor, mask):
        """encoder_output [B,T,H] fp32 CUDA, mask [B,T] bool (True = padding) or None -> [B,T]."""
        if self.training:
            raise RuntimeError("VariancePredictor: the B200 path is inference only (call .eval(); dropout is not built)")
        x = encoder_output
        if not isinstance(x, torch.Tensor) or x.dim() != 3 or x.shape[2] != self.input_size:
            raise RuntimeError(f"VariancePredictor.forward expects [B,T,{self.input_size}], got {tuple(getattr(x, 'shape', ()))}")
        if not x.is_cuda:
            raise RuntimeError("VariancePredictor.forward: input must be a CUDA tensor (this path has no CPU fallback)")
        if x.dtype != torch.float32:
            raise RuntimeError(f"VariancePredictor.forward: expected float32, got {x.dtype}")
        B, T, _ = x.shape
        if B == 0 or T == 0:
            raise RuntimeError("VariancePredictor.forward: empty batch")
        if mask is not None and tuple(mask.shape) != (B, T):
            raise RuntimeError(f"VariancePredictor.forward: mask must be [B,T] = {(B, T)}, got {tuple(mask.shape)}")
        x = x.contiguous()
        m8 = None if mask is None else mask.to(device=x.device, dtype=torch.uint8).contiguous()
        lib = _lib.lib()
        with torch.cuda.device(x.device):
            handle = self._get_handle(x.device)
            need = lib.vo2w_varpred_workspace_bytes(handle, B, T)
            if self._ws is None or self._ws.numel() < need or self._ws.device != x.device:
                self._ws = torch.empty(need, dtype=torch.uint8, device=x.device)
            out = torch.empty(B, T, dtype=torch.float32, device=x.device)
            stream = torch.cuda.current_stream(x.device).cuda_stream
            _lib.check(lib.vo2w_varpred_forward(handle, x.data_ptr(), m8.data_ptr() if m8 is not None else None, B, T,
                                                out.data_ptr(), self._ws.data_ptr(), self._ws.numel(), stream),
                       "vo2w_varpred_forward")
        return out

    def weight_blob(self) -> torch.Tensor:
        c = self.conv_layer
        parts = [c.conv1d_1.conv.weight, c.conv1d_1.conv.bias, c.layer_norm_1.weight, c.layer_norm_1.bias,
                 c.conv1d_2.conv.weight, c.conv1d_2.conv.bias, c.layer_norm_2.weight, c.layer_norm_2.bias,
                 self.linear_layer.weight, self.linear_layer.bias]
        return torch.cat([p.detach().float().cpu().reshape(-1) for p in parts]).contiguous()

    def refresh(self):
        self._drop_handle()

    def _load_from_state_dict(self, *a, **kw):
        # runs for direct and for parent-driven loads alike: the device copy of the weights is stale afterwards
        super()._load_from_state_dict(*a, **kw)
        self.refresh()

    def _get_handle(self, device):
        if self._handle is not None and self._handle_dev == device:
            return self._handle
        self._drop_handle()
        blob = self.weight_blob()
        h = C.c_void_p()
        _lib.check(_lib.lib().vo2w_varpred_create(self.input_size, self.filter_size, self.kernel, blob.data_ptr(),
                                                  blob.numel(), C.byref(h)), "vo2w_varpred_create")
        self._handle, self._handle_dev = h, device
        return h

    def _drop_handle(self):
        if getattr(self, "_handle", None) is not None:
            _lib.lib().vo2w_varpred_destroy(self._handle)
        self._handle = None

    def __del__(self):
        try:
            self._drop_handle()
        except Exception:
            pass


def variance_embed(x, value, bins, table, mean=0.0, std=1.0, control=1.0, denorm=True):
    """x [B,T,D] + table[bucketize(v, bins)] (modules.py:53-78).  denorm=True: v = ((value*std+mean)*control-mean)/std
    (the no-target branch); False: v = value (a target).  Returns (v [B,T], x + embedding)."""
    if not x.is_cuda or x.dtype != torch.float32:
        raise RuntimeError("variance_embed expects a float32 CUDA tensor (this path has no CPU fallback)")
    B, T, D = x.shape
    if tuple(value.shape) != (B, T):
        raise RuntimeError(f"variance_embed: value shape {tuple(value.shape)} does not match x {tuple(x.shape)}")
    if table.shape[0] != bins.shape[0] + 1 or table.shape[1] != D:
        raise RuntimeError("variance_embed: embedding table must be [len(bins)+1, D]")
    x = x.contiguous()
    val = value.to(device=x.device, dtype=torch.float32).contiguous()
    bins = bins.detach().to(device=x.device, dtype=torch.float32).contiguous()
    table = table.detach().to(device=x.device, dtype=torch.float32).contiguous()
    v_out = torch.empty_like(val)
    out = torch.empty_like(x)
    with torch.cuda.device(x.device):
        stream = torch.cuda.current_stream(x.device).cuda_stream
        _lib.check(_lib.lib().vo2w_variance_embed(x.data_ptr(), val.data_ptr(), bins.data_ptr(), bins.numel(), table.data_ptr(),
                                                  float(mean), float(std), float(control), int(bool(denorm)), v_out.data_ptr(),
                                                  out.data_ptr(), B * T, D, stream), "vo2w_variance_embed")
    return v_out, out


def duration_round(log_duration_prediction: torch.Tensor, d_control: float = 1.0) -> torch.Tensor:
    """clamp(round(exp(log_d) - 1) * d_control, min=0) on the device (modules.py:110-113)."""
    ld = log_duration_prediction
    if not ld.is_cuda or ld.dtype != torch.float32:
        raise RuntimeError("duration_round expects a float32 CUDA tensor (this path has no CPU fallback)")
    ld = ld.contiguous()
    out = torch.empty_like(ld)
    with torch.cuda.device(ld.device):
        stream = torch.cuda.current_stream(ld.device).cuda_stream
        _lib.check(_lib.lib().vo2w_duration_round(ld.data_ptr(), float(d_control), out.data_ptr(), ld.numel(), stream),
                   "vo2w_duration_round")
    return out


class VarianceAdaptor(nn.Module):
    """Variance Adaptor (modules.py:15-124): same constructor (reads ``stats.json`` under
    preprocess_config["path"]["preprocessed"]), same forward signature and 7-tuple result."""

    def __init__(self, preprocess_config, model_config):
        super().__init__()
        self.duration_predictor = VariancePredictor(model_config)
        self.length_regulator = LengthRegulator()
        ve = model_config["variance_embedding"]
        self.is_energy, self.is_kurtosis = ve["is_energy_condition"], ve["is_kurtosis_condition"]
        if self.is_kurtosis:
            self.kurtosis_predictor = VariancePredictor(model_config)
        if self.is_energy:
            self.energy_predictor = VariancePredictor(model_config)
        kurtosis_quantization, energy_quantization, n_bins = ve["kurtosis_quantization"], ve["energy_quantization"], ve["n_bins"]
        assert energy_quantization in ["linear", "log"]
        assert kurtosis_quantization in ["linear", "log"]
        with open(os.path.join(preprocess_config["path"]["preprocessed"], "stats.json")) as f:
            stats = json.load(f)
            energy_min, energy_max, self.energy_mean, self.energy_std = stats["energy"]
            kurt_min, kurt_max, self.kurtosis_mean, self.kurtosis_std = stats["kurtosis"]
        d_hidden = model_config["transformer"]["encoder_hidden"]

        def edges(kind, lo, hi):
            if kind == "log":
                return torch.exp(torch.linspace(np.log(lo), np.log(hi), n_bins - 1))
            return torch.linspace(lo, hi, n_bins - 1)

        self.kurt_bins = nn.Parameter(edges(kurtosis_quantization, kurt_min, kurt_max), requires_grad=False)
        self.kurt_embedding = nn.Embedding(n_bins, d_hidden)
        self.energy_bins = nn.Parameter(edges(energy_quantization, energy_min, energy_max), requires_grad=False)
        self.energy_embedding = nn.Embedding(n_bins, d_hidden)

    def _embedding(self, predictor, bins, emb, mean, std, x, target, mask, control):
        prediction = predictor(x, mask)
        if target is not None:
            _, y = variance_embed(x, target, bins, emb.weight, denorm=False)
        else:
            prediction, y = variance_embed(x, prediction, bins, emb.weight, mean, std, control, denorm=True)
        return prediction, y       # y = x + embedding (the add of modules.py:94,101 rides the same launch)

    def forward(self, x, src_mask, mel_mask=None, max_len=None, energy_target=None, kurtosis_target=None,
                duration_target=None, e_control=1.0, d_control=1.0):
        if self.training:
            raise RuntimeError("VarianceAdaptor: the B200 path is inference only (call .eval())")
        log_duration_prediction = self.duration_predictor(x, src_mask)
        energy_prediction = kurtosis_prediction = None
        if self.is_energy:
            energy_prediction, x = self._embedding(self.energy_predictor, self.energy_bins, self.energy_embedding,
                                                   self.energy_mean, self.energy_std, x, energy_target, src_mask, e_control)
        if self.is_kurtosis:
            kurtosis_prediction, x = self._embedding(self.kurtosis_predictor, self.kurt_bins, self.kurt_embedding,
                                                     self.kurtosis_mean, self.kurtosis_std, x, kurtosis_target, src_mask, 1.0)
        if duration_target is not None:
            x, mel_len = self.length_regulator(x, duration_target, max_len)
            duration_rounded = duration_target
        else:
            duration_rounded = duration_round(log_duration_prediction, d_control)
            # get_mask_from_lengths(mel_len) (modules.py:115) comes out of the same launch as the expansion
            x, mel_len, mel_mask = self.length_regulator.LR(x, duration_rounded, max_len, return_mask=True)
            if max_len:        # the reference's mask is max(mel_len) wide even when max_len crops / pads x
                mel_mask = get_mask_from_lengths(mel_len)
        return (x, energy_prediction, kurtosis_prediction, log_duration_prediction, duration_rounded, mel_len, mel_mask)
