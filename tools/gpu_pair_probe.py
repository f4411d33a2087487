"""GPU probe of the fused ResBlock-iteration kernel against a bf16-operand emulation in torch (fp64 accumulate)."""
import ctypes as C
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.nn.functional as F
from visual_onoma_to_wave_b200 import _lib

def b16(t):
    return t.to(torch.bfloat16).to(t.dtype)

def ref_iter(x, w1, b1, w2, b2, k, d):
    xa = b16(F.leaky_relu(x, 0.1)).double()    # the stream in HBM: bf16(leaky_relu(x))
    xr = torch.where(xa < 0, xa * 10.0, xa)    # the residual as the kernel recovers it
    t = F.conv1d(xa, b16(w1).double(), b1.double(), padding=d * (k - 1) // 2, dilation=d)
    t = b16(F.leaky_relu(t.float(), 0.1)).double()
    y = F.conv1d(t, b16(w2).double(), b2.double(), padding=(k - 1) // 2)
    return y + xr

def ptr_array(ts):
    return (C.c_void_p * len(ts))(*[t.data_ptr() for t in ts])

def run(Cc, ks, ds, cg, B, L, seed=0, slope=0.1):
    n = len(ks)
    g = torch.Generator().manual_seed(seed)
    w1 = [(torch.randn(Cc, Cc, k, generator=g) * (1.0 / (Cc * k) ** 0.5)).contiguous() for k in ks]
    w2 = [(torch.randn(Cc, Cc, k, generator=g) * (1.0 / (Cc * k) ** 0.5)).contiguous() for k in ks]
    b1 = [torch.randn(Cc, generator=g) * 0.1 for _ in ks]
    b2 = [torch.randn(Cc, generator=g) * 0.1 for _ in ks]
    x = [torch.randn(B, Cc, L, generator=g) for _ in ks]
    xd = [t.cuda() for t in x]
    out = torch.empty_like(xd[0])
    act = torch.empty_like(xd[0])
    scale = 1.0 / n
    ksa = (C.c_int * n)(*ks)
    dsa = (C.c_int * n)(*ds)
    lib = _lib.lib()
    t0 = time.time()
    rc = lib.vo2w_debug_pair(n, ptr_array(w1), ptr_array(b1), ptr_array(w2), ptr_array(b2), Cc, ksa, dsa, cg, ptr_array(xd), B, L,
                             scale, slope, out.data_ptr(), act.data_ptr(), None)
    _lib.check(rc, "vo2w_debug_pair")
    torch.cuda.synchronize()
    ref = sum(ref_iter(x[s], w1[s], b1[s], w2[s], b2[s], ks[s], ds[s]) for s in range(n)).float() * scale
    aref = F.leaky_relu(ref, slope)
    got = act.cpu()
    err = (got - aref).abs()
    tol = aref.abs() * 2.0 ** -7 + 2e-2   # one bf16 ulp of the value plus accumulation noise
    bad = int((err > tol).sum())
    inv = torch.where(got < 0, got / slope, got)
    aerr = (out.cpu() - inv).abs()
    abad = int((aerr > inv.abs() * 1e-6).sum())
    print(f"C={Cc:3d} k={ks} d={ds} cg={cg} B={B} L={L:6d}: max err {err.max().item():.4f} mean {err.mean().item():.5f} "
          f"bad {bad} inverse max err {aerr.max().item():.2e} bad {abad}  ({time.time() - t0:.2f}s)", flush=True)
    if bad:
        idx = (err > tol).nonzero()
        print("   first bad:", idx[:5].tolist(), "rows(l) hist:", torch.bincount((idx[:, 2] // 64).clamp(max=40))[:41].tolist())
    return bad == 0 and abad == 0

if __name__ == "__main__":
    ok = True
    cgs = [int(a) for a in sys.argv[1:]] or [1, 2]
    for cg in cgs:
        for (Cc, ks, ds, B, L) in [(32, [3], [1], 2, 700), (32, [11], [5], 3, 1500), (64, [7], [3], 2, 900), (128, [3], [1], 2, 600),
                                   (128, [11], [5], 2, 1100), (256, [3], [1], 2, 300), (256, [7], [5], 3, 500), (128, [7], [3], 5, 4000),
                                   (32, [7], [1], 1, 50), (64, [11], [1], 40, 3000),
                                   (32, [3, 7, 11], [5, 5, 5], 3, 2000), (64, [3, 7, 11], [5, 5, 5], 2, 1500),
                                   (128, [3, 7, 11], [5, 5, 5], 2, 1200), (256, [3, 7, 11], [5, 5, 5], 2, 700), (128, [3, 7], [5, 5], 2, 500)]:
            ok &= run(Cc, ks, ds, cg, B, L, slope=0.01 if len(ks) == 3 and Cc == 32 else 0.1)
    print("ALL OK" if ok else "FAILED")
    sys.exit(0 if ok else 1)
