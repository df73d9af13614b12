"""Diagnostic for the tcgen05 GEMM on a real B200: prints error structure per operand-major combination.
Usage (GPU box): python tools/gemm_diag.py > gpurun_out/gemm_diag.txt"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from m3ae_b200 import ops

torch.manual_seed(0)
dev = "cuda"

def run(M, N, K, am, bm, split=1):
    a = torch.randn((M, K) if am == 0 else (K, M), device=dev).to(torch.bfloat16)
    b = torch.randn((N, K) if bm == 0 else (K, N), device=dev).to(torch.bfloat16)
    A = a.float() if am == 0 else a.float().t()
    B = b.float() if bm == 0 else b.float().t()
    ref = A @ B.t()
    out = torch.zeros(M, N, device=dev)
    ops.gemm(a, b, out, backend=ops.GEMM_TCGEN05, a_major=am, b_major=bm, accumulate=split > 1, split_k=split)
    torch.cuda.synchronize()
    err = (out - ref).abs().max().item() / ref.abs().max().item()
    print(f"M={M} N={N} K={K} a_major={am} b_major={bm} split={split}: max_rel={err:.3e} nan={bool(torch.isnan(out).any())}")
    if err > 1e-3:
        # which K sub-chunks (of 16) inside each 64-block contribute?
        for sub in range(4):
            idx = torch.arange(K, device=dev)
            sel = ((idx % 64) // 16 == sub)
            part = A[:, sel] @ B[:, sel].t()
            print(f"   corr(out, K-subchunk {sub}) = {torch.corrcoef(torch.stack([out.flatten(), part.flatten()]))[0,1].item():.3f}")
        t = (out - ref)[:128, :256].abs()
        blk = t.view(4, 32, 4, 64).amax(dim=(1, 3))
        print("   err by (32-row, 64-col) block of tile 0:\n", blk.cpu().numpy().round(2))
        rowerr = t.amax(dim=1)
        print("   bad rows (first 32):", (rowerr > 1e-2 * ref.abs().max()).nonzero().flatten()[:32].tolist())
        colerr = t.amax(dim=0)
        print("   bad cols (first 32):", (colerr > 1e-2 * ref.abs().max()).nonzero().flatten()[:32].tolist())

for (am, bm) in [(0, 0), (0, 1), (1, 0), (1, 1)]:
    run(128, 256, 64, am, bm)
    run(128, 256, 256, am, bm)
    run(384, 768, 768, am, bm)
    run(2600, 768, 768, am, bm)
run(768, 3072, 2600, 1, 1, split=4)
print("launches", ops.launch_count())
