"""Drop-in for the acoustic model's mel ``Decoder`` (scripts/transformer/Models.py:134-197): the FFT-block stack
that LengthRegulator feeds.

Same module tree and state-dict keys as the reference (``position_enc``,
``layer_stack.{i}.slf_attn.{w_qs,w_ks,w_vs,fc,layer_norm}.*``, ``layer_stack.{i}.pos_ffn.{w_1,w_2,layer_norm}.*``),
so a vTTS checkpoint's ``decoder.*`` entries load unchanged.  ``forward`` never runs a PyTorch layer: libvo2w's
kernels do the work (csrc/decoder.cu).  Inference only: training mode (dropout, truncation to max_seq_len) is a
``RuntimeError``, like a missing CUDA device.
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn as nn

from . import _lib
from .synthetic import sinusoid_table


class _Holder(nn.Module):
    def forward(self, *a, **kw):
        raise RuntimeError(f"{type(self).__name__} is a parameter holder on the B200 path; call Decoder.forward")


class MultiHeadAttention(_Holder):
    """scripts/transformer/SubLayers.py:8-29."""

    def __init__(self, n_head, d_model, d_k, d_v, dropout=0.1):
        super().__init__()
        self.n_head, self.d_k, self.d_v = n_head, d_k, d_v
        self.w_qs = nn.Linear(d_model, n_head * d_k)
        self.w_ks = nn.Linear(d_model, n_head * d_k)
        self.w_vs = nn.Linear(d_model, n_head * d_v)
        self.layer_norm = nn.LayerNorm(d_model)
        self.fc = nn.Linear(n_head * d_v, d_model)


class PositionwiseFeedForward(_Holder):
    """scripts/transformer/SubLayers.py:65-93."""

    def __init__(self, d_in, d_hid, kernel_size, dropout=0.1):
        super().__init__()
        self.w_1 = nn.Conv1d(d_in, d_hid, kernel_size=kernel_size[0], padding=(kernel_size[0] - 1) // 2)
        self.w_2 = nn.Conv1d(d_hid, d_in, kernel_size=kernel_size[1], padding=(kernel_size[1] - 1) // 2)
        self.layer_norm = nn.LayerNorm(d_in)


class FFTBlock(_Holder):
    """scripts/transformer/Layers.py:11-19."""

    def __init__(self, d_model, n_head, d_k, d_v, d_inner, kernel_size, dropout=0.1):
        super().__init__()
        self.slf_attn = MultiHeadAttention(n_head, d_model, d_k, d_v, dropout=dropout)
        self.pos_ffn = PositionwiseFeedForward(d_model, d_inner, kernel_size, dropout=dropout)


class Decoder(nn.Module):
    """Decoder(config): ``config`` is the model.yaml dict (max_seq_len, transformer.decoder_*, conv_filter_size,
    conv_kernel_size), as in the reference."""

    def __init__(self, config):
        super().__init__()
        t = config["transformer"]
        d_model, n_head, n_layers = t["decoder_hidden"], t["decoder_head"], t["decoder_layer"]
        d_k = d_v = d_model // n_head
        self.max_seq_len = config["max_seq_len"]
        self.d_model, self.n_head, self.d_inner = d_model, n_head, t["conv_filter_size"]
        self.kernel_size = list(t["conv_kernel_size"])
        self.position_enc = nn.Parameter(sinusoid_table(self.max_seq_len + 1, d_model).unsqueeze(0), requires_grad=False)
        self.layer_stack = nn.ModuleList([
            FFTBlock(d_model, n_head, d_k, d_v, self.d_inner, self.kernel_size, dropout=t.get("decoder_dropout", 0.1))
            for _ in range(n_layers)])
        self._handle = None
        self._handle_dev = None
        self._ws = None
        self._pos = None
        self._long = None

    # ---- reference surface -------------------------------------------------------------------
    def forward(self, enc_seq: torch.Tensor, mask: torch.Tensor, return_attns: bool = False):
        """enc_seq [B,L,d_model] fp32, mask [B,L] bool (True = padding frame) -> (dec_output [B,L,d_model], mask)."""
        if self.training:
            raise RuntimeError("Decoder: the B200 path is inference only (call .eval(); dropout is not built)")
        if return_attns:
            raise RuntimeError("Decoder.forward: attention maps are never materialised on the B200 path (return_attns=True)")
        if not isinstance(enc_seq, torch.Tensor) or enc_seq.dim() != 3 or enc_seq.shape[2] != self.d_model:
            raise RuntimeError(f"Decoder.forward expects [B,L,{self.d_model}], got {tuple(getattr(enc_seq, 'shape', ()))}")
        if not enc_seq.is_cuda:
            raise RuntimeError("Decoder.forward: input must be a CUDA tensor (this path has no CPU fallback)")
        if enc_seq.dtype != torch.float32:
            raise RuntimeError(f"Decoder.forward: expected float32, got {enc_seq.dtype}")
        B, L, D = enc_seq.shape
        if B == 0 or L == 0:
            raise RuntimeError("Decoder.forward: empty batch")
        if mask is None or tuple(mask.shape) != (B, L):
            raise RuntimeError(f"Decoder.forward: mask must be [B,L] = {(B, L)}, got {tuple(getattr(mask, 'shape', ()))}")
        x = enc_seq.contiguous()
        m8 = mask.to(device=x.device, dtype=torch.uint8).contiguous()
        lib = _lib.lib()
        with torch.cuda.device(x.device):
            handle = self._get_handle(x.device)
            pos = self._pos_rows(L, x.device)
            need = lib.vo2w_decoder_workspace_bytes(handle, B, L)
            if self._ws is None or self._ws.numel() < need or self._ws.device != x.device:
                self._ws = None
                self._ws = torch.empty(need, dtype=torch.uint8, device=x.device)
            out = torch.empty_like(x)
            stream = torch.cuda.current_stream(x.device).cuda_stream
            _lib.check(lib.vo2w_decoder_forward(handle, x.data_ptr(), pos.data_ptr(), m8.data_ptr(), B, L, out.data_ptr(),
                                                self._ws.data_ptr(), self._ws.numel(), stream), "vo2w_decoder_forward")
        return out, mask

    # ---- plumbing ----------------------------------------------------------------------------
    def _pos_rows(self, L, device):
        """First L rows of the position table; a longer table is built the reference's way when L > max_seq_len
        (Models.py:175-181)."""
        if L <= self.max_seq_len:
            if self._pos is None or self._pos.device != device or self._pos.shape[0] != self.max_seq_len + 1:
                self._pos = self.position_enc.detach()[0].to(device=device, dtype=torch.float32).contiguous()
            return self._pos
        if self._long is None or self._long.device != device or self._long.shape[0] < L:
            self._long = sinusoid_table(L, self.d_model).to(device)
        return self._long

    def weight_blob(self) -> torch.Tensor:
        """fp32 [n] in vo2w_decoder_create's order."""
        parts = []
        for blk in self.layer_stack:
            a, f = blk.slf_attn, blk.pos_ffn
            parts += [a.w_qs.weight, a.w_ks.weight, a.w_vs.weight, a.w_qs.bias, a.w_ks.bias, a.w_vs.bias,
                      a.fc.weight, a.fc.bias, a.layer_norm.weight, a.layer_norm.bias,
                      f.w_1.weight, f.w_1.bias, f.w_2.weight, f.w_2.bias, f.layer_norm.weight, f.layer_norm.bias]
        return torch.cat([p.detach().float().cpu().reshape(-1) for p in parts]).contiguous()

    def refresh(self):
        """Re-upload the weights on the next forward (after load_state_dict / in-place edits)."""
        self._drop_handle()
        self._pos = None

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
        _lib.check(_lib.lib().vo2w_decoder_create(self.d_model, self.n_head, self.d_inner, self.kernel_size[0],
                                                  self.kernel_size[1], len(self.layer_stack), blob.data_ptr(), blob.numel(),
                                                  C.byref(h)), "vo2w_decoder_create")
        self._handle, self._handle_dev = h, device
        return h

    def _drop_handle(self):
        if getattr(self, "_handle", None) is not None:
            _lib.lib().vo2w_decoder_destroy(self._handle)
        self._handle = None

    def __del__(self):
        try:
            self._drop_handle()
        except Exception:
            pass
