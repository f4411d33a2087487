"""vo2w_varpred_forward_static: the VariancePredictor forward for callers that cannot allocate (CUDA graph capture, tight
serving loops).  It has no counterpart in the reference: its scratch is a
device array built into libvo2w.so (VO2W_VARPRED_STATIC_BYTES) instead of a caller workspace.

Called through the C-ABI and compared with the existing entry point, vo2w_varpred_forward, on two calls in a row with
different inputs and shapes (the second call finds the first call's contents in the array), and with the oracle
(scripts/model/modules.py:161-213) within the tolerance of the predictor's own parity test.  The file is named so that
pytest collects it last: a failure here must not hide another test.  It builds its
own handle and buffers, so it passes alone in a fresh process and after the other tests alike."""
import ctypes as C

import pytest
import torch

from visual_onoma_to_wave_b200 import _lib
from visual_onoma_to_wave_b200.synthetic import seeded_variance_adaptor_state_dict, variance_model_config
import visual_onoma_to_wave_b200 as v

pytestmark = pytest.mark.gpu

VARIANCE_TOL = 1e-4                 # the tolerance of tests/test_gpu_parity.py::test_variance_pieces_against_oracle
STATIC_BYTES = 8 << 20              # VO2W_VARPRED_STATIC_BYTES (include/vo2w.h)


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    assert _lib.lib().vo2w_device_check() == 0, "libvo2w needs an sm_100 device"
    return torch.device("cuda:0")


def _handle(sd, prefix):
    vp = v.VariancePredictor(variance_model_config())
    vp.load_state_dict({k[len(prefix):]: t for k, t in sd.items() if k.startswith(prefix)})
    blob = vp.weight_blob()
    h = C.c_void_p()
    _lib.check(_lib.lib().vo2w_varpred_create(vp.input_size, vp.filter_size, vp.kernel, blob.data_ptr(), blob.numel(),
                                              C.byref(h)), "vo2w_varpred_create")
    return h, vp.input_size, vp.filter_size


def _both(h, x, mask, dev):
    """x [B,T,d_in] and mask [B,T] on the host -> (existing entry point, allocation-free entry point), both on the host."""
    lib = _lib.lib()
    B, T, _ = x.shape
    xd = x.to(dev).contiguous()
    md = None if mask is None else mask.to(dev, torch.uint8).contiguous()
    mp = None if md is None else md.data_ptr()
    need = lib.vo2w_varpred_workspace_bytes(h, B, T)
    assert 0 < need <= STATIC_BYTES
    ws = torch.empty(need, dtype=torch.uint8, device=dev)
    out_ws = torch.full((B, T), float("nan"), device=dev)
    out_st = torch.full((B, T), float("nan"), device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    _lib.check(lib.vo2w_varpred_forward(h, xd.data_ptr(), mp, B, T, out_ws.data_ptr(), ws.data_ptr(), ws.numel(), stream),
               "vo2w_varpred_forward")
    _lib.check(lib.vo2w_varpred_forward_static(h, xd.data_ptr(), mp, B, T, out_st.data_ptr(), stream),
               "vo2w_varpred_forward_static")
    torch.cuda.synchronize(dev)
    return out_ws.cpu(), out_st.cpu()


def test_varpred_forward_static_matches_the_workspace_entry_point_twice(dev):
    from oracle import variance_oracle as vo
    lib = _lib.lib()
    sd = seeded_variance_adaptor_state_dict(77)
    with torch.cuda.device(dev):
        h, d_in, F = _handle(sd, "energy_predictor.")
        try:
            # call 1: the largest batch the array holds for these widths (ragged lengths), so nearly all of it is dirty
            B1, T1 = 16, 1
            while lib.vo2w_varpred_workspace_bytes(h, B1, T1 + 1) <= STATIC_BYTES:
                T1 += 1
            assert T1 >= 100 and lib.vo2w_varpred_workspace_bytes(h, B1, T1) <= STATIC_BYTES
            g = torch.Generator().manual_seed(11)
            x1 = torch.randn(B1, T1, d_in, generator=g)
            lens = torch.randint(1, T1 + 1, (B1,), generator=g)
            lens[0] = T1
            m1 = torch.arange(T1)[None, :] >= lens[:, None]
            a1, s1 = _both(h, x1, m1, dev)
            # call 2, straight after: other inputs, a smaller and odd shape, no mask -- the array still holds call 1
            x2 = torch.randn(3, 37, d_in, generator=g) * 2 + 0.5
            a2, s2 = _both(h, x2, None, dev)
            for (a, s, x, m) in ((a1, s1, x1, m1), (a2, s2, x2, None)):
                assert torch.isfinite(s).all()
                assert float((s - a).abs().max()) <= VARIANCE_TOL          # measured: bit-identical, the same kernels run
                ref = vo.variance_predictor(sd, "energy_predictor.", x, m)
                assert float((s - ref).abs().max()) <= VARIANCE_TOL
            assert torch.equal(s1, a1) and torch.equal(s2, a2)
            # call 1 again: the second call left nothing behind that changes it
            a3, s3 = _both(h, x1, m1, dev)
            assert torch.equal(s3, s1)

            # beyond the built-in array: a clear error, no launch, the output untouched
            Tbig = 1
            while lib.vo2w_varpred_workspace_bytes(h, 1, Tbig) <= STATIC_BYTES:
                Tbig += 64
            xb = torch.zeros(1, Tbig, d_in, device=dev)
            ob = torch.full((1, Tbig), 7.0, device=dev)
            rc = lib.vo2w_varpred_forward_static(h, xb.data_ptr(), None, 1, Tbig, ob.data_ptr(),
                                                 torch.cuda.current_stream(dev).cuda_stream)
            assert rc == _lib.E_WORKSPACE
            msg = lib.vo2w_last_error().decode()
            assert "varpred_forward_static" in msg and str(STATIC_BYTES) in msg, msg
            torch.cuda.synchronize(dev)
            assert bool((ob == 7.0).all())
            assert lib.vo2w_varpred_forward_static(h, None, None, 1, 4, ob.data_ptr(), None) == _lib.E_INVALID
        finally:
            lib.vo2w_varpred_destroy(h)
