"""Host logic of the tcgen05 path, checked on CPU: weight packing, tap tables and the
epilogue's index mapping are replayed in numpy exactly as conv_umma.cu walks them and
compared with F.conv1d / F.conv_transpose1d.  (The MMA itself is checked on the GPU.)"""
import ctypes as C

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from visual_onoma_to_wave_b200 import _lib

PAD, TILE = 32, 128


def padded_len(L):
    return PAD + (L + TILE - 1) // TILE * TILE + PAD


def bf16_round(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(torch.bfloat16).float().numpy()


def pack(transposed, w, b, Cin, Cout, k, ds):
    lib = _lib.lib()
    meta = np.zeros(12 + 5 * 12, dtype=np.int32)
    w = np.ascontiguousarray(w, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    _lib.check(lib.vo2w_debug_pack(int(transposed), w.ctypes.data, b.ctypes.data, Cin, Cout, k, ds, meta.ctypes.data,
                                   None, 0, None, 0))
    wb = np.zeros(meta[10], dtype=np.uint16)
    bb = np.zeros(meta[11], dtype=np.float32)
    _lib.check(lib.vo2w_debug_pack(int(transposed), w.ctypes.data, b.ctypes.data, Cin, Cout, k, ds, meta.ctypes.data,
                                   wb.ctypes.data, wb.size, bb.ctypes.data, bb.size))
    return meta, wb, bb


def replay_kernel(meta, wb, bb, x):
    """x: [B,Cin,L] float32 (already activated).  Mirrors producer / MMA / epilogue indexing of conv_umma.cu."""
    Cin_p, Cout, n_cols, n_tiles, ppt, halo_l, halo_r, n_items, w_tile_stride, up_s = [int(v) for v in meta[:10]]
    items = meta[12:12 + 5 * n_items].reshape(n_items, 5)
    B, Cin, L = x.shape
    Lp = padded_len(L)
    xin = np.zeros((B, Cin_p // 8, Lp, 8), dtype=np.float32)
    xr = bf16_round(x)
    for c in range(Cin):
        xin[:, c // 8, PAD:PAD + L, c % 8] = xr[:, c, :]
    wf = (wb.astype(np.uint32) << 16).view(np.float32)
    rows = TILE + halo_l + halo_r
    Cout_p = (Cout + 15) // 16 * 16 if up_s == 1 else Cout
    out = np.zeros((B, Cout_p, L * up_s), dtype=np.float64)
    for b in range(B):
        for mt in range((L + TILE - 1) // TILE):
            l0 = mt * TILE
            start = PAD + l0 - halo_l
            assert start >= 0 and start + rows <= Lp
            A = xin[b, :, start:start + rows, :]                       # [chunk][row][8] as staged in smem
            A2 = A.transpose(1, 0, 2).reshape(rows, Cin_p).astype(np.float64)
            for ty in range(n_tiles):
                D = np.full((TILE, n_cols), np.nan)
                extra = (ty * ppt + up_s // 2) // up_s if up_s > 1 else 0
                for (row_shift, n_off, n_size, zero_init, w_off) in items:
                    base = (ty * w_tile_stride + w_off) // 2
                    Wt = wf[base:base + n_size * Cin_p].reshape(Cin_p // 8, n_size, 8)
                    Wn = Wt.transpose(1, 0, 2).reshape(n_size, Cin_p).astype(np.float64)
                    assert row_shift + extra + TILE <= rows
                    prod = A2[row_shift + extra:row_shift + extra + TILE] @ Wn.T
                    if zero_init:
                        D[:, n_off:n_off + n_size] = prod
                    else:
                        D[:, n_off:n_off + n_size] += prod
                assert not np.isnan(D).any()
                for q in range(TILE):
                    row = l0 + q
                    if row >= L:
                        continue
                    for n in range(n_cols):
                        if up_s == 1:
                            co, lo = ty * n_cols + n, row
                        else:
                            ph = n // Cout
                            co, lo = n - ph * Cout, up_s * row + ty * ppt + ph
                        out[b, co, lo] = D[q, n] + bb[ty * n_cols + n]
    return out[:, :Cout]


CONV_CASES = [(32, 32, 3, 1), (32, 32, 11, 5), (64, 64, 7, 3), (80, 512, 7, 1), (32, 1, 7, 1), (256, 256, 3, 5)]


@pytest.mark.parametrize("Cin,Cout,k,d", CONV_CASES)
def test_conv_packing_replay(Cin, Cout, k, d):
    g = torch.Generator().manual_seed(Cin + k + d)
    w = torch.randn(Cout, Cin, k, generator=g) * 0.1
    b = torch.randn(Cout, generator=g)
    x = torch.randn(2, Cin, 150, generator=g)
    meta, wb, bb = pack(False, w.numpy(), b.numpy(), Cin, Cout, k, d)
    got = replay_kernel(meta, wb, bb, x.numpy())
    ref = F.conv1d(x.to(torch.bfloat16).double(), w.to(torch.bfloat16).double(), b.double(), padding=d * (k - 1) // 2,
                   dilation=d).numpy()
    np.testing.assert_allclose(got, ref, atol=1e-9)


@pytest.mark.parametrize("Cin,Cout,s", [(512, 256, 8), (256, 128, 8), (128, 64, 2), (64, 32, 2), (32, 16, 4)])
def test_convT_packing_replay(Cin, Cout, s):
    g = torch.Generator().manual_seed(Cin + s)
    k = 2 * s
    w = torch.randn(Cin, Cout, k, generator=g) * 0.1
    b = torch.randn(Cout, generator=g)
    x = torch.randn(1, Cin, 131, generator=g)
    meta, wb, bb = pack(True, w.numpy(), b.numpy(), Cin, Cout, k, s)
    got = replay_kernel(meta, wb, bb, x.numpy())
    ref = F.conv_transpose1d(x.to(torch.bfloat16).double(), w.to(torch.bfloat16).double(), b.double(), stride=s,
                             padding=s // 2).numpy()
    np.testing.assert_allclose(got, ref, atol=1e-9)


def test_pack_rejects_unsupported_shapes():
    w = np.zeros((16, 16, 4), dtype=np.float32)
    b = np.zeros(16, dtype=np.float32)
    meta = np.zeros(72, dtype=np.int32)
    lib = _lib.lib()
    rc = lib.vo2w_debug_pack(0, w.ctypes.data, b.ctypes.data, 16, 16, 4, 1, meta.ctypes.data, None, 0, None, 0)
    assert rc == _lib.E_INVALID and b"kernel size" in lib.vo2w_last_error()
    with pytest.raises(RuntimeError):
        _lib.check(lib.vo2w_debug_pack(1, w.ctypes.data, b.ctypes.data, 16, 16, 4, 3, meta.ctypes.data, None, 0, None, 0))
