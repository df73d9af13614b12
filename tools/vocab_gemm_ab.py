"""Times the decoder's vocabulary projection (M=128, N=30522, K=768, bias + arg-max partials) on the tcgen05 GEMM with wide (128x256)
and skinny (128x64) tiles, L2 flushed between launches.  Usage (GPU box): python tools/vocab_gemm_ab.py"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import m3ae_b200
from m3ae_b200 import ops
dev="cuda"
B, D, V = 128, 768, 30522
x = torch.randn(B, D, device=dev).to(torch.bfloat16)
w = (0.02*torch.randn(V, D, device=dev)).to(torch.bfloat16)
bias = torch.zeros(V, device=dev)
big = torch.empty(256*1024*1024//4, device=dev)  # L2 flush buffer
def run(mode):
    ops.set_gemm_skinny(mode)
    n_grp = (V + 63)//64
    av, ai = torch.empty(n_grp, B, device=dev), torch.empty(n_grp, B, dtype=torch.int32, device=dev)
    def f(): ops.gemm_group([ops.GemmSpec(x, w, None, bias=bias, amax=(av, ai))])
    for _ in range(3): f()
    torch.cuda.synchronize()
    ts=[]
    for _ in range(10):
        big.zero_()  # evict w from L2 (as the 141 MB KV stream does in the decoder step)
        e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        e0.record(); f(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1)*1e3)
    ts.sort(); print("mode", mode, "median us", round(ts[len(ts)//2],1), "min", round(ts[0],1))
for m in (0,1,2): run(m)
ops.set_gemm_skinny(0)
