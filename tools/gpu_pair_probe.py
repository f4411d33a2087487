"""GPU probe of the fused ResBlock-iteration kernel against a bf16-operand emulation in torch (fp64 accumulate)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.nn.functional as F
from visual_onoma_to_wave_b200 import _lib

def b16(t):
    return t.to(torch.bfloat16).to(t.dtype)

def ref_pair(x, w1, b1, w2, b2, k, d):
    xd = b16(x).double()                       # stream value in HBM is bf16
    xa = b16(F.leaky_relu(xd, 0.1))
    t = F.conv1d(xa, b16(w1).double(), b1.double(), padding=d * (k - 1) // 2, dilation=d)
    t = b16(F.leaky_relu(t.float(), 0.1).double())
    y = F.conv1d(t, b16(w2).double(), b2.double(), padding=(k - 1) // 2)
    return (y + xd).float()

def run(C, k, d, cg, B, L, seed=0):
    g = torch.Generator().manual_seed(seed)
    w1 = torch.randn(C, C, k, generator=g) * (1.0 / (C * k) ** 0.5)
    w2 = torch.randn(C, C, k, generator=g) * (1.0 / (C * k) ** 0.5)
    b1 = torch.randn(C, generator=g) * 0.1
    b2 = torch.randn(C, generator=g) * 0.1
    x = torch.randn(B, C, L, generator=g)
    xd = x.cuda()
    out = torch.empty_like(xd)
    act = torch.empty_like(xd)
    lib = _lib.lib()
    t0 = time.time()
    rc = lib.vo2w_debug_pair(w1.contiguous().data_ptr(), b1.data_ptr(), w2.contiguous().data_ptr(), b2.data_ptr(), C, k, d, cg,
                             xd.data_ptr(), B, L, out.data_ptr(), act.data_ptr(), None)
    _lib.check(rc, "vo2w_debug_pair")
    torch.cuda.synchronize()
    ref = ref_pair(x, w1, b1, w2, b2, k, d)
    got = out.cpu()
    err = (got - ref).abs()
    # out is rounded to bf16: allow one bf16 ulp of the value plus accumulation noise
    tol = ref.abs() * 2.0 ** -7 + 2e-2
    bad = int((err > tol).sum())
    aerr = (act.cpu() - b16(F.leaky_relu(got, 0.1))).abs().max().item()
    print(f"C={C:3d} k={k:2d} d={d} cg={cg} B={B} L={L:6d}: max err {err.max().item():.4f} mean {err.mean().item():.5f} "
          f"bad {bad} act-consistency {aerr:.2e}  ({time.time() - t0:.2f}s)", flush=True)
    if bad:
        idx = (err > tol).nonzero()
        print("   first bad:", idx[:5].tolist(), "rows(l) hist:", torch.bincount((idx[:, 2] // 64).clamp(max=40))[:41].tolist())
    return bad == 0 and aerr == 0.0

if __name__ == "__main__":
    ok = True
    cgs = [int(a) for a in sys.argv[1:]] or [1, 2]
    for cg in cgs:
        for (C, k, d, B, L) in [(32, 3, 1, 2, 700), (32, 11, 5, 3, 1500), (64, 7, 3, 2, 900), (128, 3, 1, 2, 600), (128, 11, 5, 2, 1100),
                                (256, 3, 1, 2, 300), (256, 7, 5, 3, 500), (128, 7, 3, 5, 4000), (32, 7, 1, 1, 50), (64, 11, 1, 40, 3000)]:
            ok &= run(C, k, d, cg, B, L)
    print("ALL OK" if ok else "FAILED")
    sys.exit(0 if ok else 1)
