"""Prints where a small tcgen05 GEMM disagrees with torch (layout debugging aid)."""
import sys
import torch
sys.path.insert(0, ".")
from deepfm_b200.ops.gemm import gemm_bf16

def probe(M, N, K, a_mn, b_mn):
    torch.manual_seed(0)
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    a_in = a.t().contiguous() if a_mn else a
    b_in = b.t().contiguous() if b_mn else b
    c = gemm_bf16(a_in, b_in, a_mn_major=a_mn, b_mn_major=b_mn)
    torch.cuda.synchronize()
    ref = a.float() @ b.float().t()
    err = (c - ref).abs()
    ok = err <= (2e-2 + 2e-3 * ref.abs())
    print(f"M{M} N{N} K{K} a_mn={a_mn} b_mn={b_mn}: max_err={err.max().item():.4f} "
          f"bad={(~ok).sum().item()}/{ok.numel()}")
    if (~ok).any():
        rows_bad = (~ok).any(1).nonzero().flatten()[:16].tolist()
        cols_bad = (~ok).any(0).nonzero().flatten()[:16].tolist()
        print("  first bad rows", rows_bad, "cols", cols_bad)
        print("  c[0,:8]  ", c[0, :8].tolist())
        print("  ref[0,:8]", ref[0, :8].tolist())
        # does it match a GEMM over a prefix of K?  (descriptor K-advance bugs)
        for kk in (16, 32, 48, 64):
            if kk <= K:
                p = a[:, :kk].float() @ b[:, :kk].float().t()
                print(f"  match with K prefix {kk}: {(c - p).abs().max().item():.4f}")

for a_mn, b_mn in [(False, False), (True, False), (False, True), (True, True)]:
    for shape in [(128, 64, 64), (128, 64, 16), (128, 256, 128), (256, 128, 256)]:
        try:
            probe(*shape, a_mn, b_mn)
        except Exception as e:
            print("EXC", shape, a_mn, b_mn, repr(e)[:300])
            sys.exit(1)
