"""Host logic of stream_umma_kernel (conv_pre and the upsamplers of the bf16 mode), checked on CPU: the weight packing
per CTA rank, the item / tile shift tables and the epilogue's (phase, channel) mapping are replayed in numpy exactly
as stream_umma.cu walks them -- producer rows, the N/2 split of the B operand between the two CTAs of a pair, row
blocks, N-tiles -- and compared with F.conv1d / F.conv_transpose1d (scripts/hifigan/models.py:118-135).  The MMA
itself is checked on the GPU."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from visual_onoma_to_wave_b200 import _lib

PAD, TILE = 32, 128
MAX_ITEMS, MAX_TILES = 12, 16


def padded_len(L):
    return PAD + (L + TILE - 1) // TILE * TILE + PAD


def bf16_round(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(torch.bfloat16).float().numpy()


def pack(transposed, w, b, Cin, Cout, k, ds):
    lib = _lib.lib()
    meta = np.zeros(16 + MAX_ITEMS + MAX_TILES, dtype=np.int32)
    w = np.ascontiguousarray(w, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    _lib.check(lib.vo2w_debug_stream_pack(int(transposed), w.ctypes.data, b.ctypes.data, Cin, Cout, k, ds, meta.ctypes.data,
                                          None, 0, None, 0))
    wb = np.zeros(meta[12], dtype=np.uint16)
    bb = np.zeros(meta[13], dtype=np.float32)
    _lib.check(lib.vo2w_debug_stream_pack(int(transposed), w.ctypes.data, b.ctypes.data, Cin, Cout, k, ds, meta.ctypes.data,
                                          wb.ctypes.data, wb.size, bb.ctypes.data, bb.size))
    return meta, wb, bb


def replay_kernel(meta, wb, bb, x, Cout_true):
    """x: [B,Cin,L] float32 (already activated).  Mirrors stream_umma.cu: input producer, issuer, epilogue."""
    Cin_p, Cout, up_s, KB, n_kb, NB, n_cols, n_tiles, n_items, halo_l, halo_r, stage_bytes = [int(v) for v in meta[:12]]
    ph_tile, cb = int(meta[14]), int(meta[15])
    cblks = Cout // cb if up_s > 1 else n_tiles
    assert ph_tile * cb == n_cols and (up_s == 1 or (Cout % cb == 0 and n_tiles == (up_s // ph_tile) * cblks))
    item_shift = meta[16:16 + MAX_ITEMS]
    tile_shift = meta[16 + MAX_ITEMS:16 + MAX_ITEMS + MAX_TILES]
    n_half, kb8 = n_cols // 2, KB // 8
    assert stage_bytes == n_half * KB * 2 and 2 * NB * n_cols <= 512 and n_cols % 32 == 0
    B, Cin, L = x.shape
    Lp = padded_len(L)
    xin = np.zeros((B, Cin_p // 8, Lp, 8), dtype=np.float64)
    xr = bf16_round(x)
    for c in range(Cin):
        xin[:, c // 8, PAD:PAD + L, c % 8] = xr[:, c, :]
    wf = (wb.astype(np.uint32) << 16).view(np.float32).astype(np.float64)
    a_rows = TILE * NB + halo_l + halo_r
    stage = n_half * KB
    out = np.full((B, Cout, L * up_s), np.nan)
    for b in range(B):
        for mt in range((L + TILE * NB - 1) // (TILE * NB)):
            l0 = mt * TILE * NB
            start = PAD + l0 - halo_l
            rows = min(a_rows, Lp - start)
            assert start >= 0 and rows > 0
            A = np.full((Cin_p // 8, a_rows, 8), np.nan)               # rows the short last tile does not stage stay NaN
            A[:, :rows] = xin[b, :, start:start + rows]
            for nt in range(n_tiles):
                D = np.zeros((NB, TILE, n_cols))
                for i in range(n_items):
                    row = int(item_shift[i] + tile_shift[nt])
                    assert 0 <= row and row + TILE * NB <= a_rows
                    for kb in range(n_kb):
                        for rank in range(2):
                            s = (((rank * n_tiles + nt) * n_items + i) * n_kb + kb) * stage
                            Wn = wf[s:s + stage].reshape(kb8, n_half, 8).transpose(1, 0, 2).reshape(n_half, KB)
                            for blk in range(NB):
                                Ar = A[kb * kb8:(kb + 1) * kb8, blk * TILE + row:blk * TILE + row + TILE].transpose(1, 0, 2).reshape(TILE, KB)
                                live = min(TILE, max(0, L - l0 - blk * TILE))     # rows >= L are never stored; they may read NaN
                                D[blk, :live, rank * n_half:(rank + 1) * n_half] += Ar[:live] @ Wn.T
                for blk in range(NB):
                    for q in range(TILE):
                        r = l0 + blk * TILE + q
                        if r >= L:
                            continue
                        n = np.arange(n_cols)                             # column -> (phase, channel): phase-major inside a tile
                        ph, co = (nt // cblks) * ph_tile + n // cb, (nt % cblks) * cb + n % cb
                        out[b, co, up_s * r + ph] = D[blk, q] + bb[nt * n_cols + n]
    assert not np.isnan(out).any()
    return out[:, :Cout_true]


@pytest.mark.parametrize("Cin,Cout,k,d", [(80, 512, 7, 1), (32, 32, 3, 1), (64, 64, 7, 3), (32, 1, 7, 1), (256, 256, 3, 5)])
def test_stream_conv_packing_replay(Cin, Cout, k, d):
    g = torch.Generator().manual_seed(Cin + k + d)
    w = torch.randn(Cout, Cin, k, generator=g) * 0.1
    b = torch.randn(Cout, generator=g)
    x = torch.randn(2, Cin, 300, generator=g)
    meta, wb, bb = pack(False, w.numpy(), b.numpy(), Cin, Cout, k, d)
    got = replay_kernel(meta, wb, bb, x.numpy(), Cout)
    ref = F.conv1d(x.to(torch.bfloat16).double(), w.to(torch.bfloat16).double(), b.double(), padding=d * (k - 1) // 2,
                   dilation=d).numpy()
    np.testing.assert_allclose(got, ref, atol=1e-9)


@pytest.mark.parametrize("Cin,Cout,k,s", [(512, 256, 16, 8), (256, 128, 16, 8), (128, 64, 4, 2), (64, 32, 4, 2), (32, 32, 8, 4),
                                           (64, 64, 8, 2)])
def test_stream_convT_packing_replay(Cin, Cout, k, s):
    g = torch.Generator().manual_seed(Cin + s)
    w = torch.randn(Cin, Cout, k, generator=g) * 0.1
    b = torch.randn(Cout, generator=g)
    x = torch.randn(1, Cin, 300, generator=g)
    meta, wb, bb = pack(True, w.numpy(), b.numpy(), Cin, Cout, k, s)
    got = replay_kernel(meta, wb, bb, x.numpy(), Cout)
    ref = F.conv_transpose1d(x.to(torch.bfloat16).double(), w.to(torch.bfloat16).double(), b.double(), stride=s,
                             padding=(k - s) // 2).numpy()
    np.testing.assert_allclose(got, ref, atol=1e-9)


def test_stream_geometry_of_the_generator_layers():
    """The tile shapes DESIGN.md section 4.3 states for the V1 Generator: NB x n_cols, weights resident for s = 2."""
    exp = {(1, 512, 256, 16, 8): (1, 256, 8, 2), (1, 256, 128, 16, 8): (2, 128, 8, 2), (1, 128, 64, 4, 2): (2, 128, 1, 3),
           (1, 64, 32, 4, 2): (4, 64, 1, 3), (0, 80, 512, 7, 1): (2, 128, 4, 7)}
    for (tr, Cin, Cout, k, ds), (NB, n_cols, n_tiles, n_items) in exp.items():
        w = np.zeros((Cin, Cout, k) if tr else (Cout, Cin, k), dtype=np.float32)
        meta, _, _ = pack(tr, w, np.zeros(Cout, dtype=np.float32), Cin, Cout, k, ds)
        assert (int(meta[5]), int(meta[6]), int(meta[7]), int(meta[8])) == (NB, n_cols, n_tiles, n_items), (tr, Cin, Cout)


def test_stream_pack_rejects_unsupported_shapes():
    lib = _lib.lib()
    meta = np.zeros(16 + MAX_ITEMS + MAX_TILES, dtype=np.int32)
    w = np.zeros((16, 16, 4), dtype=np.float32)
    b = np.zeros(16, dtype=np.float32)
    assert lib.vo2w_debug_stream_pack(0, w.ctypes.data, b.ctypes.data, 16, 16, 4, 1, meta.ctypes.data, None, 0, None, 0) == _lib.E_INVALID
    assert lib.vo2w_debug_stream_pack(1, w.ctypes.data, b.ctypes.data, 16, 16, 4, 2, meta.ctypes.data, None, 0, None, 0) == _lib.E_INVALID
