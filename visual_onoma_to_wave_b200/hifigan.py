"""Drop-in for the reference's ``hifigan`` package (scripts/hifigan/__init__.py, models.py).

``Generator(h)`` keeps the reference's module tree and state-dict names
(``conv_pre``, ``ups.{i}``, ``resblocks.{n}.convs{1,2}.{m}``, ``conv_post`` with
``{bias, weight_g, weight_v}`` before and ``{weight, bias}`` after
``remove_weight_norm()``), so ``get_vocoder`` (scripts/utils/model.py:56-68) works
unchanged: construct, ``load_state_dict(ckpt["generator"])``, ``eval()``,
``remove_weight_norm()``, ``to(device)``.  ``forward`` does not run PyTorch modules: it
hands the folded weights to libvo2w.so once and then calls the sm_100a kernels through the
C ABI (include/vo2w.h).  No CUDA device or no library => RuntimeError, never a fallback.
"""
from __future__ import annotations

import ctypes as C
import math
import os
from collections import OrderedDict

import torch
import torch.nn as nn

from . import _lib

LRELU_SLOPE = 0.1   # scripts/hifigan/models.py:7


class AttrDict(dict):
    """scripts/hifigan/__init__.py:4-7."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.__dict__ = self


def get_padding(kernel_size, dilation=1):
    """scripts/hifigan/models.py:16-17."""
    return int((kernel_size * dilation - dilation) / 2)


class _WNConv(nn.Module):
    """Parameter holder for one weight-normalised Conv1d / ConvTranspose1d.

    Same parameter names as ``weight_norm(nn.Conv1d(...))`` in the reference
    (models.py:26-92,118-145): ``bias, weight_g, weight_v``; after ``fold()`` ``weight, bias``.
    Default init = nn.Conv1d's (the reference's ``init_weights`` has no lasting effect,
    SURVEY.md section 3.2); g = ||v|| over all dims but 0.
    """

    def __init__(self, shape, n_bias, kernel_size, stride=1, dilation=1, padding=0, transposed=False):
        super().__init__()
        self.kernel_size, self.stride, self.dilation, self.padding = kernel_size, stride, dilation, padding
        self.transposed = transposed
        v = torch.empty(shape)
        nn.init.kaiming_uniform_(v, a=math.sqrt(5))
        bound = 1.0 / math.sqrt(shape[1] * shape[2])
        self.bias = nn.Parameter(torch.empty(n_bias).uniform_(-bound, bound))
        self.weight_g = nn.Parameter(v.reshape(shape[0], -1).norm(dim=1).reshape(shape[0], 1, 1))
        self.weight_v = nn.Parameter(v)

    @property
    def folded(self) -> bool:
        return "weight" in self._parameters

    def folded_weight(self) -> torch.Tensor:
        if self.folded:
            return self.weight.detach()
        # what remove_weight_norm() leaves behind; dim 0 is Cin for ConvTranspose1d
        return torch._weight_norm(self.weight_v.detach(), self.weight_g.detach(), 0)

    def fold(self):
        if self.folded:
            raise ValueError("weight_norm of 'weight' not found in {}".format(self))   # torch's message
        w = self.folded_weight()
        del self._parameters["weight_g"]
        del self._parameters["weight_v"]
        self.weight = nn.Parameter(w)     # appended after bias, as torch's remove_weight_norm does


class ResBlock(nn.Module):
    """Parameter tree of scripts/hifigan/models.py:20-109 (ResBlock1-type; the only one the fork has)."""

    def __init__(self, h, channels, kernel_size=3, dilation=(1, 3, 5)):
        super().__init__()
        self.h = h
        self.convs1 = nn.ModuleList(
            _WNConv((channels, channels, kernel_size), channels, kernel_size, dilation=d,
                    padding=get_padding(kernel_size, d)) for d in dilation)
        self.convs2 = nn.ModuleList(
            _WNConv((channels, channels, kernel_size), channels, kernel_size, dilation=1,
                    padding=get_padding(kernel_size, 1)) for _ in dilation)

    def remove_weight_norm(self):
        for l in list(self.convs1) + list(self.convs2):
            l.fold()


_COPY_POOL = None


def _host_copy(dst: torch.Tensor, src: torch.Tensor, workers: int = 4) -> None:
    """dst.copy_(src) over row chunks on a few threads of our own (copy_ releases the GIL).  torchrun exports
    OMP_NUM_THREADS=1 to every rank, which makes a plain copy_ of a 180 MB result single-threaded and costs the
    host-buffer path ~10 % (the last sub-batch's copy is not hidden behind kernels)."""
    global _COPY_POOL
    n = dst.shape[0]
    if n < 2 * workers or dst.numel() * dst.element_size() < (8 << 20):
        dst.copy_(src)
        return
    if _COPY_POOL is None:
        from concurrent.futures import ThreadPoolExecutor
        _COPY_POOL = ThreadPoolExecutor(max_workers=workers, thread_name_prefix="vo2w-copy")
    step = (n + workers - 1) // workers
    futs = [_COPY_POOL.submit(dst[i:i + step].copy_, src[i:i + step]) for i in range(0, n, step)]
    for f in futs:
        f.result()


def pipeline_chunks(B: int, sub: int):
    """(start, end) of the sub-batches one host call walks.  Full sub-batches in the middle; the first one and the last
    ones are smaller (down to a quarter, from sub-batches of 64 up): what the two-stream pipeline cannot hide is the
    host->device copy in front of the first forward and the device->host copy behind the last one, and both are
    proportional to their sub-batch (at 8 ranks on one host the last copy of 23 MB per rank was 3 % of the step).  A
    batch-shape change costs one small launch (zero_pad_rows), and a forward is linear in its batch down to 32."""
    if B <= sub or sub < 64:
        return [(a, min(a + sub, B)) for a in range(0, B, sub)]
    small = sub // 4
    sizes = [small]
    left = B - small
    while left > sub:
        sizes.append(sub)
        left -= sub
    while left > 2 * small:                         # taper: half of what is left, ..., then two pieces of about `small`
        sizes.append(left // 2)
        left -= left // 2
    if left > small:
        sizes.append(left - left // 2)
        left //= 2
    if left:
        sizes.append(left)
    out, a = [], 0
    for n in sizes:
        out.append((a, a + n))
        a += n
    assert a == B and all(0 < e - b <= sub for b, e in out)
    return out


class Generator(nn.Module):
    """HiFi-GAN generator, mel [B,80,T] -> wav [B,1,T*prod(upsample_rates)] (models.py:112-174).

    precision: "bf16" (tcgen05 tensor-core kernels, SNR >= 40 dB vs the fp32 reference) or
    "fp32" (tcgen05 kind::tf32 kernels with a hi/lo operand split, max-abs-err <= 1e-4).  Default from $VO2W_PRECISION, else "bf16".
    """

    def __init__(self, h, precision: str | None = None):
        super().__init__()
        self.h = h
        self.num_kernels = len(h.resblock_kernel_sizes)
        self.num_upsamples = len(h.upsample_rates)
        self.precision = precision or os.environ.get("VO2W_PRECISION", "bf16")
        if self.precision not in ("bf16", "fp32"):
            raise ValueError(f"precision must be 'bf16' or 'fp32', got {self.precision!r}")
        c0 = h.upsample_initial_channel
        self.conv_pre = _WNConv((c0, 80, 7), c0, 7, padding=3)        # 80 is hard-coded (models.py:119)
        self.ups = nn.ModuleList()
        for i, (u, k) in enumerate(zip(h.upsample_rates, h.upsample_kernel_sizes)):
            self.ups.append(_WNConv((c0 // 2 ** i, c0 // 2 ** (i + 1), k), c0 // 2 ** (i + 1), k, stride=u,
                                    padding=(k - u) // 2, transposed=True))
        self.resblocks = nn.ModuleList()
        ch = c0
        for i in range(len(self.ups)):
            ch = c0 // 2 ** (i + 1)
            for k, d in zip(h.resblock_kernel_sizes, h.resblock_dilation_sizes):
                self.resblocks.append(ResBlock(h, ch, k, d))
        self.conv_post = _WNConv((1, ch, 7), 1, 7, padding=3)
        self._handle = None
        self._handle_key = None
        self._ws = OrderedDict()        # (B,T) -> workspace tensor (zero halo rows are an invariant of ours)
        self._host = {}                 # (B,T) -> pinned staging for the host-buffer call
        self.last_launches = 0
        self.host_sub_batch = 64
        self._pipe = None

    # ---- reference surface -------------------------------------------------------------
    def remove_weight_norm(self):
        print("Removing weight norm...")          # models.py:168
        for l in self.ups:
            l.fold()
        for l in self.resblocks:
            l.remove_weight_norm()
        self.conv_pre.fold()
        self.conv_post.fold()
        self._drop_handle()

    def load_state_dict(self, *args, **kwargs):
        out = super().load_state_dict(*args, **kwargs)
        self._drop_handle()
        return out

    def _apply(self, fn, *args, **kwargs):
        out = super()._apply(fn, *args, **kwargs)
        self._drop_handle()
        return out

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not isinstance(x, torch.Tensor) or x.dim() != 3 or x.shape[1] != 80:
            raise RuntimeError(f"Generator.forward expects mel [B,80,T], got {tuple(getattr(x, 'shape', ()))}")
        if not x.is_cuda:
            raise RuntimeError("Generator.forward: input must be a CUDA tensor (this path has no CPU fallback)")
        if x.dtype != torch.float32:
            raise RuntimeError(f"Generator.forward: expected float32 mel, got {x.dtype}")
        B, _, T = x.shape
        if B == 0 or T == 0:
            raise RuntimeError("Generator.forward: empty batch")
        x = x.contiguous()
        lib = _lib.lib()
        with torch.cuda.device(x.device):
            handle = self._get_handle(x.device)
            ws, fresh = self._workspace(handle, B, T, x.device)
            wav = torch.empty(B, 1, T * self.hop, dtype=torch.float32, device=x.device)
            stream = torch.cuda.current_stream(x.device).cuda_stream
            _lib.check(lib.vo2w_generator_forward(handle, x.data_ptr(), B, T, wav.data_ptr(), ws.data_ptr(),
                                                  ws.numel(), int(fresh), stream), "vo2w_generator_forward")
            self.last_launches = lib.vo2w_generator_last_launches(handle)
        return wav

    # ---- beyond the reference: host-buffer call used by vocoder_infer and bench e2e -------
    @property
    def hop(self) -> int:
        return int(math.prod(self.h.upsample_rates))

    def infer_host(self, mel_host: torch.Tensor, device=None, int16_scale: float | None = None,
                   sub_batch: int | None = None, out: torch.Tensor | None = None):
        """Host mel [B,80,T] -> host wav through the C ABI's host-buffer entry points.

        Up to ``sub_batch`` (default ``self.host_sub_batch`` = 64) utterances: one vo2w_generator_forward_host call
        (H2D, kernels, D2H, synchronise) into a pinned buffer owned by the generator.  Larger batches are walked in
        sub-batches on two alternating CUDA streams (vo2w_generator_forward_host_async), each with its own
        workspace and pinned staging, so the copies of one sub-batch overlap the kernels of the other; the result is
        a fresh host tensor.  Returns float32 [B, T*hop], or int16 when ``int16_scale`` (max_wav_value) is given.

        ``out``: optional caller-owned host tensor [B, T*hop] of the result dtype, ideally pinned (a serving loop keeps
        one per shard): every device->host copy then lands in it directly -- no staging copy, no allocation, no
        first-touch page faults -- and it is returned.
        """
        if mel_host.is_cuda or mel_host.dtype != torch.float32 or mel_host.dim() != 3 or mel_host.shape[1] != 80:
            raise RuntimeError("infer_host expects a CPU float32 mel [B,80,T]")
        device = torch.device(device if device is not None else self._device())
        if device.type != "cuda":
            raise RuntimeError("infer_host: the vocoder must live on a CUDA device (no CPU fallback)")
        B, _, T = mel_host.shape
        mel_host = mel_host.contiguous()
        sub = int(sub_batch or self.host_sub_batch)
        if out is not None:
            odt = torch.int16 if int16_scale is not None else torch.float32
            if out.is_cuda or out.dtype != odt or tuple(out.shape) != (B, T * self.hop) or not out.is_contiguous():
                raise RuntimeError(f"infer_host: out must be a contiguous CPU {odt} tensor of shape {(B, T * self.hop)}")
        # False: the result aliases a pinned buffer of the generator that the next call reuses
        self.host_out_owned = B > sub or out is not None
        if B > sub:
            return self._infer_host_pipelined(mel_host, device, int16_scale, sub, out)
        lib = _lib.lib()
        with torch.cuda.device(device):
            handle = self._get_handle(device)
            ws, fresh = self._workspace(handle, B, T, device)
            key = (B, T, device.index, int16_scale is not None)
            if key not in self._host:
                self._host.clear()
                odt = torch.int16 if int16_scale is not None else torch.float32
                self._host[key] = (torch.empty(B, 80, T, device=device), torch.empty(B, T * self.hop, device=device),
                                   torch.empty(B, T * self.hop, dtype=odt).pin_memory())
            mel_dev, wav_dev, pin = self._host[key]
            out = pin if out is None else out
            stream = torch.cuda.current_stream(device).cuda_stream
            f32_ptr = None if int16_scale is not None else out.data_ptr()
            i16_ptr = out.data_ptr() if int16_scale is not None else None
            _lib.check(lib.vo2w_generator_forward_host(handle, mel_host.data_ptr(), B, T, f32_ptr, i16_ptr,
                                                       float(int16_scale or 0.0), ws.data_ptr(), ws.numel(), int(fresh),
                                                       mel_dev.data_ptr(), wav_dev.data_ptr(), stream),
                       "vo2w_generator_forward_host")
            self.last_launches = lib.vo2w_generator_last_launches(handle)
        return out

    def _infer_host_pipelined(self, mel_host, device, int16_scale, sub, out=None):
        lib = _lib.lib()
        B, _, T = mel_host.shape
        odt = torch.int16 if int16_scale is not None else torch.float32
        n_out = T * self.hop
        direct = out is not None                  # copy device -> caller's buffer, no staging
        if out is None:
            out = torch.empty(B, n_out, dtype=odt)
        with torch.cuda.device(device):
            handle = self._get_handle(device)
            key = (sub, T, device.index, odt)
            if self._pipe is None or self._pipe[0] != key:
                need = lib.vo2w_generator_workspace_bytes(handle, sub, T)
                slots = []
                for _ in range(2):
                    slots.append({"stream": torch.cuda.Stream(device), "ws": torch.empty(need, dtype=torch.uint8, device=device),
                                  "mel": torch.empty(sub, 80, T, device=device), "wav": torch.empty(sub, n_out, device=device),
                                  "pin": torch.empty(sub, n_out, dtype=odt).pin_memory(), "shape": None, "pending": None})
                self._pipe = (key, slots)
            slots = self._pipe[1]
            cur = torch.cuda.current_stream(device)
            for sl in slots:
                sl["stream"].wait_stream(cur)
            launches = 0

            def retire(sl):
                if sl["pending"] is not None:
                    a, b = sl["pending"]
                    sl["stream"].synchronize()
                    if not direct:
                        _host_copy(out[a:b], sl["pin"][: b - a])  # host copy; the other stream's kernels keep running
                    sl["pending"] = None

            for i, (a, b) in enumerate(pipeline_chunks(B, sub)):
                sl = slots[i & 1]
                # Results copied straight into the caller's buffer need no host step between sub-batches: a slot's
                # workspace / staging tensors are reused in stream order, so the whole call is queued at once and the GPU
                # never waits for the host to notice a finished sub-batch (both slots used to finish together, and the
                # next H2D + 34 launches then started from an idle GPU).  With pinned staging the host copy comes first.
                if not direct:
                    retire(sl)
                torch.cuda.nvtx.range_push(f"vo2w.sub_batch[{a}:{b}]")
                fresh = sl["shape"] != (b - a, T)            # the workspace layout depends on the batch size
                sl["shape"] = (b - a, T)
                src = mel_host[a:b]
                dst = out[a:b] if direct else sl["pin"]
                f32_ptr = None if int16_scale is not None else dst.data_ptr()
                i16_ptr = dst.data_ptr() if int16_scale is not None else None
                _lib.check(lib.vo2w_generator_forward_host_async(handle, src.data_ptr(), b - a, T, f32_ptr, i16_ptr,
                                                                 float(int16_scale or 0.0), sl["ws"].data_ptr(), sl["ws"].numel(),
                                                                 int(fresh), sl["mel"].data_ptr(), sl["wav"].data_ptr(),
                                                                 sl["stream"].cuda_stream),
                           "vo2w_generator_forward_host_async")
                launches += lib.vo2w_generator_last_launches(handle)
                sl["pending"] = (a, b)
                torch.cuda.nvtx.range_pop()
            for sl in slots:
                retire(sl)
            self.last_launches = launches
        return out

    def debug_tap(self, name: str) -> torch.Tensor:
        """fp32 activation of the last forward ("ups.0", and with VO2W_DEBUG_TAPS=1 "conv_pre", "mrf.i")."""
        if self._handle is None:
            raise RuntimeError("debug_tap: no forward has run")
        buf = torch.empty(64 * 1024 * 1024, dtype=torch.float32)
        n = _lib.lib().vo2w_generator_debug_tap(self._handle, name.encode(), buf.data_ptr(), buf.numel())
        _lib.check(int(n), "vo2w_generator_debug_tap")
        return buf[:n].clone()

    def has_tap(self, name: str) -> bool:
        """Whether the last forward left the activation `name` in the workspace (a tensor a fused launch never writes
        to HBM -- e.g. "mrf_act.3" when conv_post rides the last MRF launch -- has no tap)."""
        if self._handle is None:
            return False
        return bool(_lib.lib().vo2w_generator_has_tap(self._handle, name.encode()))

    def launch_times(self, x: torch.Tensor):
        """One forward with a CUDA event around every launch: ([ms per launch], [kind per launch]), kind 0 = layout
        helper, 1 = single convolution, 2 = fused ResBlock iterations (bench.py's roofline measurement)."""
        import ctypes as C
        lib = _lib.lib()
        handle = self._get_handle(x.device)
        _lib.check(lib.vo2w_generator_set_timing(handle, 1), "vo2w_generator_set_timing")
        try:
            self.forward(x)
            ms = (C.c_float * 256)()
            kind = (C.c_int32 * 256)()
            n = lib.vo2w_generator_get_timing(handle, ms, kind, 256)
            _lib.check(n, "vo2w_generator_get_timing")
        finally:
            lib.vo2w_generator_set_timing(handle, 0)
        return list(ms[:n]), list(kind[:n])

    def refresh(self):
        """Re-pack the weights on the next forward (call after editing parameters in place)."""
        self._drop_handle()

    # ---- plumbing ------------------------------------------------------------------------
    def _device(self):
        return self.conv_pre.bias.device

    def _layers(self):
        yield self.conv_pre
        yield from self.ups
        for rb in self.resblocks:
            yield from rb.convs1
            yield from rb.convs2
        yield self.conv_post

    def folded_blob(self) -> torch.Tensor:
        """fp32 [n] = for every layer weight then bias, in state-dict order (the ABI's weights_host)."""
        parts = []
        for l in self._layers():
            parts.append(l.folded_weight().float().cpu().reshape(-1))
            parts.append(l.bias.detach().float().cpu().reshape(-1))
        return torch.cat(parts).contiguous()

    def _c_config(self) -> _lib.GenConfig:
        h = self.h
        cfg = _lib.GenConfig()
        cfg.num_upsamples = len(h.upsample_rates)
        cfg.num_kernels = len(h.resblock_kernel_sizes)
        if cfg.num_upsamples > 4 or cfg.num_kernels > 3:
            raise RuntimeError("libvo2w supports at most 4 upsampling stages and 3 resblock kernels")
        for i, (u, k) in enumerate(zip(h.upsample_rates, h.upsample_kernel_sizes)):
            cfg.upsample_rates[i], cfg.upsample_kernel_sizes[i] = u, k
        cfg.upsample_initial_channel = h.upsample_initial_channel
        for j, (k, d) in enumerate(zip(h.resblock_kernel_sizes, h.resblock_dilation_sizes)):
            cfg.resblock_kernel_sizes[j] = k
            for m in range(3):
                cfg.resblock_dilation_sizes[j][m] = d[m]
        cfg.num_mels = 80
        return cfg

    def _drop_handle(self):
        if getattr(self, "_handle", None) is not None:
            _lib.lib().vo2w_generator_destroy(self._handle)
        self._handle = None
        self._handle_key = None
        if hasattr(self, "_ws"):
            self._ws.clear()
            self._host.clear()
            self._pipe = None

    def _get_handle(self, device):
        key = (device.index if device.index is not None else torch.cuda.current_device(), self.precision)
        if self._handle is not None and key == self._handle_key:
            return self._handle
        self._drop_handle()
        blob = self.folded_blob()
        cfg = self._c_config()
        handle = C.c_void_p()
        prec = _lib.PREC_BF16 if self.precision == "bf16" else _lib.PREC_FP32
        _lib.check(_lib.lib().vo2w_generator_create(C.byref(cfg), blob.data_ptr(), blob.numel(), prec, C.byref(handle)),
                   "vo2w_generator_create")
        self._handle, self._handle_key = handle, key
        return handle

    def _workspace(self, handle, B, T, device):
        # one workspace per (shape, device, stream): two forwards of one shape on different streams must not share it
        key = (B, T, device.index, torch.cuda.current_stream(device).cuda_stream)
        if key in self._ws:
            self._ws.move_to_end(key)
            return self._ws[key], False
        need = _lib.lib().vo2w_generator_workspace_bytes(handle, B, T)
        while len(self._ws) >= 2:
            self._ws.popitem(last=False)
        self._ws[key] = torch.empty(need, dtype=torch.uint8, device=device)
        return self._ws[key], True

    def __del__(self):
        try:
            self._drop_handle()
        except Exception:
            pass
