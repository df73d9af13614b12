"""Times LayerNorm fwd/bwd at the fusion stack's sizes (bf16, H=768): backward with the bulk-copy staging ring vs the
register-prefetch path, with / without hidden dropout (cached keep bits or regenerated Philox).
Usage (GPU box): python tools/ln_bench.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from m3ae_b200 import ops

dev, bf, H = "cuda", torch.bfloat16, 768


def timeit(fn, reps=30):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


for M in (20800, 2048):
    # several buffers, rotated, so that consecutive launches do not hit in the 126 MB L2
    nb = 6 if M > 4096 else 1
    xs = [torch.randn(M, H, device=dev).to(bf) for _ in range(nb)]
    dys = [torch.randn(M, H, device=dev).to(bf) for _ in range(nb)]
    gamma, beta = torch.randn(H, device=dev), torch.randn(H, device=dev)
    y, mean, rstd = ops.layernorm_fwd(xs[0], gamma, beta, 1e-12)
    bits = torch.randint(-2**31, 2**31 - 1, (M, H // 32), dtype=torch.int32, device=dev)
    dx, dxd = torch.empty_like(xs[0]), torch.empty_like(xs[0])
    dg, db, dc = (torch.zeros(H, device=dev) for _ in range(3))
    it = [0]

    def fwd():
        it[0] += 1
        ops.layernorm_fwd(xs[it[0] % nb], gamma, beta, 1e-12, y, mean, rstd)

    us = timeit(fwd)
    print(f"M={M:6d} ln_fwd                                   {us:7.1f} us  {2 * M * H * 2 / us / 1e3:7.1f} GB/s (x in, y out)")
    for regs_only in (True, False):
        ops.set_layernorm_bwd_staging(regs_only)
        for name, kw, nbytes in (("no dropout", dict(), 3), ("dropout, cached bits", dict(drop_p=0.1, drop_seed=1, drop_stream=1, drop_bits=bits), 4),
                                 ("dropout, Philox regenerated", dict(drop_p=0.1, drop_seed=1, drop_stream=1), 4)):
            def bwd():
                it[0] += 1
                i = it[0] % nb
                ops.L.check(ops.L.load().m3ae_layernorm_bwd(
                    dys[i].data_ptr(), xs[i].data_ptr(), mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr(), dx.data_ptr(), dg.data_ptr(),
                    db.data_ptr(), dc.data_ptr(), M, H, ops.BF16, dxd.data_ptr() if kw else None, kw.get("drop_p", 0.0), kw.get("drop_seed", 0),
                    kw.get("drop_stream", 0), kw["drop_bits"].data_ptr() if "drop_bits" in kw else None, ops._stream()), "ln_bwd")
            us = timeit(bwd)
            print(f"M={M:6d} ln_bwd {'registers' if regs_only else 'ring     '} {name:28s} {us:7.1f} us  {nbytes * M * H * 2 / us / 1e3:7.1f} GB/s")
    ops.set_layernorm_bwd_staging(False)
