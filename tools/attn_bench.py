"""Times the attention kernels at the fusion-stack shapes (B=64, 12 heads x 64). Usage: python tools/attn_bench.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from m3ae_b200 import ops
dev = "cuda"
B, H, dh = int(os.environ.get("B", "64")), 12, 64
reps = int(os.environ.get("REPS", "10"))
ops.set_attention_backend(int(os.environ.get("ATTN_BACKEND", "0")))  # 0 auto (tcgen05), 3 mma.sync
def run(Lq, Lk, drop=0.0):
    q = torch.randn(B * Lq, 3 * H * dh, device=dev).to(torch.bfloat16)
    kv = torch.randn(B * Lk, 3 * H * dh, device=dev).to(torch.bfloat16)
    Q, K, V = q[:, :768], kv[:, 768:1536], kv[:, 1536:]
    out = torch.empty(B * Lq, 768, device=dev, dtype=torch.bfloat16)
    lse = torch.empty(B, H, Lq, device=dev)
    dout = torch.randn_like(out)
    dq = torch.empty_like(q); dkv = torch.empty_like(kv)
    delta = torch.empty(B, H, Lq, device=dev)
    kw = dict(B=B, H=H, Lq=Lq, Lk=Lk, dh=dh, drop_p=drop, drop_seed=1, drop_stream=0)
    pregen = os.environ.get("PREGEN", "0") == "1" and drop > 0
    cache = os.environ.get("CACHE", "0") == "1" and drop > 0
    if pregen or cache:  # production: keep bits generated ahead by m3ae_dropout_bits / cached by the forward, read by the backward
        kw["drop_mask"] = torch.zeros(B * H * Lq * ops.drop_mask_ld(Lk), dtype=torch.uint8, device=dev)
    if pregen:
        gen = ops.DropoutBits([(0, kw["drop_mask"], B * H * Lq, Lk, 0, drop, 0)])
        gen.run([1])
        kw["drop_mask_ready"] = True
        def g(): gen.run([1])
    def f(): ops.attn_fwd(Q, K, V, out, lse, **kw)
    def b(): ops.attn_bwd(Q, K, V, out, lse, dout=dout, dq=dq[:, :768], dk=dkv[:, 768:1536], dv=dkv[:, 1536:], delta=delta, **kw)
    for name, fn, mult in (("fwd", f, 4), ("bwd", b, 10)) + ((("bitgen", g, 0),) if pregen else ()):
        for _ in range(2): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if os.environ.get("GRAPH", "1") == "1":
            # a ctypes call costs ~15 us of host time, more than the small kernels run: replay `reps` launches from a CUDA graph
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                fn()
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr, stream=side):
                    for _ in range(reps): fn()
            torch.cuda.current_stream().wait_stream(side)
            gr.replay(); torch.cuda.synchronize()
            e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
        else:
            e0.record()
            for _ in range(reps): fn()
            e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / reps
        flops = mult * B * H * Lq * Lk * dh
        print(f"attn {name} Lq={Lq:4d} Lk={Lk:4d} drop={drop}: {us:8.1f} us  {flops / us / 1e6:7.1f} TFLOP/s (algorithmic {mult}*B*H*Lq*Lk*dh)")
shapes = [tuple(int(v) for v in t.split("x")) for t in os.environ.get("SHAPES", "325x325,325x32,32x325,32x32").split(",")]
drops = [float(v) for v in os.environ.get("DROPS", "0.0,0.1").split(",")]
for (Lq, Lk) in shapes:
    for drop in drops:
        run(Lq, Lk, drop)
