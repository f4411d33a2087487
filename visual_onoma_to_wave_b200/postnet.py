"""Drop-in for the acoustic model's tail: ``PostNet`` (scripts/transformer/Layers.py:32-137) and the
``mel_linear`` projection in front of it (scripts/model/vtts.py:49-52,106-107).

Same module tree and state-dict keys as the reference (``convolutions.{i}.0.conv.{weight,bias}``,
``convolutions.{i}.1.{weight,bias,running_mean,running_var,num_batches_tracked}``), so a vTTS checkpoint's
``postnet.*`` entries load unchanged.  ``forward`` never runs a PyTorch layer: the eval-mode BatchNorm is folded
into each convolution (as weight_norm is for the vocoder) and libvo2w's kernels do the work.  Inference only:
training mode (dropout p=0.5, batch statistics) is a ``RuntimeError``, like a missing CUDA device.
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn as nn

from . import _lib


class ConvNorm(nn.Module):
    """scripts/transformer/Layers.py:32-64: a Conv1d holder (``.conv``; this fork never applies ``w_init_gain``)."""

    def __init__(self, in_channels, out_channels, kernel_size=1, stride=1, padding=None, dilation=1, bias=True,
                 w_init_gain="linear"):
        super().__init__()
        if padding is None:
            assert kernel_size % 2 == 1
            padding = int(dilation * (kernel_size - 1) / 2)
        self.conv = nn.Conv1d(in_channels, out_channels, kernel_size=kernel_size, stride=stride, padding=padding,
                              dilation=dilation, bias=bias)

    def forward(self, signal):
        raise RuntimeError("ConvNorm is a parameter holder on the B200 path; call PostNet.forward")


class PostNet(nn.Module):
    """PostNet: five 1-d convolutions with 512 channels and kernel size 5 (Layers.py:67-137)."""

    def __init__(self, n_mel_channels=80, postnet_embedding_dim=512, postnet_kernel_size=5, postnet_n_convolutions=5):
        super().__init__()
        self.n_mel, self.dim, self.k, self.n_conv = n_mel_channels, postnet_embedding_dim, postnet_kernel_size, postnet_n_convolutions
        pad = int((postnet_kernel_size - 1) / 2)
        self.convolutions = nn.ModuleList()
        chans = [n_mel_channels] + [postnet_embedding_dim] * (postnet_n_convolutions - 1) + [n_mel_channels]
        for i in range(postnet_n_convolutions):
            self.convolutions.append(nn.Sequential(
                ConvNorm(chans[i], chans[i + 1], kernel_size=postnet_kernel_size, stride=1, padding=pad, dilation=1,
                         w_init_gain="tanh" if i < postnet_n_convolutions - 1 else "linear"),
                nn.BatchNorm1d(chans[i + 1])))
        self._handle = None
        self._handle_dev = None
        self._ws = None

    # ---- reference surface -------------------------------------------------------------------
    def forward(self, x: torch.Tensor, add_input: bool = False) -> torch.Tensor:
        """x [B,L,n_mel] -> [B,L,n_mel].  ``add_input=True`` returns ``postnet(x) + x`` (vtts.py:107) in the same pass."""
        if self.training:
            raise RuntimeError("PostNet: the B200 path is inference only (call .eval(); dropout / batch statistics are not built)")
        if not isinstance(x, torch.Tensor) or x.dim() != 3 or x.shape[2] != self.n_mel:
            raise RuntimeError(f"PostNet.forward expects [B,L,{self.n_mel}], got {tuple(getattr(x, 'shape', ()))}")
        if not x.is_cuda:
            raise RuntimeError("PostNet.forward: input must be a CUDA tensor (this path has no CPU fallback)")
        if x.dtype != torch.float32:
            raise RuntimeError(f"PostNet.forward: expected float32, got {x.dtype}")
        B, L, _ = x.shape
        if B == 0 or L == 0:
            raise RuntimeError("PostNet.forward: empty batch")
        x = x.contiguous()
        lib = _lib.lib()
        with torch.cuda.device(x.device):
            handle = self._get_handle(x.device)
            need = lib.vo2w_postnet_workspace_bytes(handle, B, L)
            if self._ws is None or self._ws.numel() < need or self._ws.device != x.device:
                self._ws = torch.empty(need, dtype=torch.uint8, device=x.device)
            out = torch.empty_like(x)
            stream = torch.cuda.current_stream(x.device).cuda_stream
            _lib.check(lib.vo2w_postnet_forward(handle, x.data_ptr(), B, L, out.data_ptr(), int(bool(add_input)),
                                                self._ws.data_ptr(), self._ws.numel(), stream), "vo2w_postnet_forward")
        return out

    # ---- plumbing ----------------------------------------------------------------------------
    def folded_blob(self) -> torch.Tensor:
        """fp32 [n]: per convolution weight [Cout,Cin,k] then bias [Cout], eval-mode BatchNorm folded in."""
        parts = []
        for seq in self.convolutions:
            conv, bn = seq[0].conv, seq[1]
            w, b = conv.weight.detach().double().cpu(), conv.bias.detach().double().cpu()
            g = bn.weight.detach().double().cpu() / torch.sqrt(bn.running_var.detach().double().cpu() + bn.eps)
            parts.append((w * g[:, None, None]).float().reshape(-1))
            parts.append(((b - bn.running_mean.detach().double().cpu()) * g + bn.bias.detach().double().cpu()).float())
        return torch.cat(parts).contiguous()

    def refresh(self):
        """Re-fold the weights on the next forward (after load_state_dict / in-place edits)."""
        self._drop_handle()

    def _load_from_state_dict(self, *a, **kw):
        # runs for direct and for parent-driven loads alike: the device copy of the weights is stale afterwards
        super()._load_from_state_dict(*a, **kw)
        self.refresh()

    def _get_handle(self, device):
        if self._handle is not None and self._handle_dev == device:
            return self._handle
        self._drop_handle()
        blob = self.folded_blob()
        h = C.c_void_p()
        _lib.check(_lib.lib().vo2w_postnet_create(self.n_mel, self.dim, self.k, self.n_conv, blob.data_ptr(), blob.numel(),
                                                  C.byref(h)), "vo2w_postnet_create")
        self._handle, self._handle_dev = h, device
        return h

    def _drop_handle(self):
        if getattr(self, "_handle", None) is not None:
            _lib.lib().vo2w_postnet_destroy(self._handle)
        self._handle = None

    def __del__(self):
        try:
            self._drop_handle()
        except Exception:
            pass


def mel_linear(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor | None = None) -> torch.Tensor:
    """``nn.Linear`` on the last axis through libvo2w (vtts.py:106 ``self.mel_linear(output)``): x [...,K] fp32 CUDA."""
    if not x.is_cuda or x.dtype != torch.float32:
        raise RuntimeError("mel_linear expects a float32 CUDA tensor (this path has no CPU fallback)")
    N, K = weight.shape
    if x.shape[-1] != K:
        raise RuntimeError(f"mel_linear: input feature size {x.shape[-1]} does not match weight {tuple(weight.shape)}")
    x2 = x.contiguous().reshape(-1, K)
    w = weight.detach().to(device=x.device, dtype=torch.float32).contiguous()
    b = None if bias is None else bias.detach().to(device=x.device, dtype=torch.float32).contiguous()
    y = torch.empty(x2.shape[0], N, dtype=torch.float32, device=x.device)
    if x2.shape[0]:
        with torch.cuda.device(x.device):
            scratch = torch.empty(2 * N * K, dtype=torch.float32, device=x.device)     # packed weights (tensor-pipe path)
            stream = torch.cuda.current_stream(x.device).cuda_stream
            _lib.check(_lib.lib().vo2w_linear_forward(x2.data_ptr(), w.data_ptr(), b.data_ptr() if b is not None else None,
                                                      x2.shape[0], K, N, y.data_ptr(), scratch.data_ptr(), stream),
                       "vo2w_linear_forward")
    return y.reshape(*x.shape[:-1], N)
