"""Times every GEMM stage of one fusion layer pair (B=64: text M=2048, image M=20800) on the tcgen05 kernel.
Usage (GPU box): python tools/gemm_bench.py > gpurun_out/gemm_bench.txt"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from m3ae_b200 import ops
from m3ae_b200.functional import _wgrad_split

dev = "cuda"
B = int(os.environ.get("B", "64"))
Mt, Mi, H, FF = B * 32, B * 325, 768, 3072
bf = torch.bfloat16

def t(*s, dtype=bf):
    return (torch.randn(*s, device=dev) * 0.05).to(dtype)

ONLY = os.environ.get("ONLY", "")
ops.set_gemm_cta_pair(int(os.environ.get("PAIR", "0")))  # 0 auto (two-pair clusters with multicast), 1 single-CTA tiles, 2 pairs, 3 two-pair clusters
print(f"# m3ae_set_gemm_cta_pair({os.environ.get('PAIR', '0')})")
REPS = int(os.environ.get("REPS", "20"))

def bench(name, mk_specs, reps=None):
    if ONLY and ONLY not in name:
        return
    reps = reps or REPS
    specs = mk_specs()
    flops = sum(2.0 * s.problem.M * s.problem.N * s.problem.K for s in specs)
    for _ in range(3):
        ops.gemm_group(specs)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        ops.gemm_group(specs)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / reps
    print(f"{name:46s} {us:9.1f} us  {flops / us / 1e6:8.1f} TFLOP/s   units={[ (s.problem.M, s.problem.N, s.problem.K, s.problem.split_k) for s in specs]}")

MN = ops.MAJOR_MN
xs = {"t": t(Mt, H), "i": t(Mi, H)}
ff = {"t": t(Mt, FF), "i": t(Mi, FF)}
wqkv, wo, w1, w2, wkv = t(3 * H, H), t(H, H), t(FF, H), t(H, FF), t(2 * H, H)
bias = {n: torch.randn(n, device=dev) for n in (H, 2 * H, 3 * H, FF)}
def out(M, N, dtype=bf): return torch.empty(M, N, device=dev, dtype=dtype)
Ms = {"t": Mt, "i": Mi}

for which in (["i"], ["t"], ["t", "i"]):
    tag = "+".join(which)
    bench(f"[{tag}] fwd QKV N=2304 bias", lambda: [ops.GemmSpec(xs[s], wqkv, out(Ms[s], 3 * H), bias=bias[3 * H]) for s in which])
    bench(f"[{tag}] fwd O-proj N=768 bias+res", lambda: [ops.GemmSpec(xs[s], wo, out(Ms[s], H), bias=bias[H], residual=xs[s]) for s in which])
    bench(f"[{tag}] fwd O-proj N=768 bias+drop+res", lambda: [ops.GemmSpec(xs[s], wo, out(Ms[s], H), bias=bias[H], residual=xs[s], drop_p=0.1, drop_seed=1, drop_stream=1) for s in which])
    bench(f"[{tag}] fwd plain N=768 no epilogue", lambda: [ops.GemmSpec(xs[s], wo, out(Ms[s], H)) for s in which])
    bench(f"[{tag}] fwd FFN-up N=3072 bias+gelu+aux", lambda: [ops.GemmSpec(xs[s], w1, out(Ms[s], FF), bias=bias[FF], act=ops.ACT_GELU, aux_out=out(Ms[s], FF)) for s in which])
    bench(f"[{tag}] fwd FFN-up N=3072 bias+gelu+aux=gelu'", lambda: [ops.GemmSpec(xs[s], w1, out(Ms[s], FF), bias=bias[FF], act=ops.ACT_GELU, aux_out=out(Ms[s], FF), aux_is_grad=True) for s in which])
    bench(f"[{tag}] fwd FFN-up N=3072 plain", lambda: [ops.GemmSpec(xs[s], w1, out(Ms[s], FF)) for s in which])
    bench(f"[{tag}] fwd FFN-down K=3072 bias+res", lambda: [ops.GemmSpec(ff[s], w2, out(Ms[s], H), bias=bias[H], residual=xs[s]) for s in which])
    bench(f"[{tag}] dgrad FFN-down (B MN) *gelu'", lambda: [ops.GemmSpec(xs[s], w2, out(Ms[s], FF), b_major=MN, aux_in=ff[s], dact=ops.ACT_GELU) for s in which])
    bench(f"[{tag}] dgrad FFN-down (B MN) *aux", lambda: [ops.GemmSpec(xs[s], w2, out(Ms[s], FF), b_major=MN, aux_in=ff[s], dact=ops.ACT_GELU, aux_is_grad=True) for s in which])
    bench(f"[{tag}] dgrad FFN-up (B MN) +res", lambda: [ops.GemmSpec(ff[s], w1, out(Ms[s], H), b_major=MN, residual=xs[s]) for s in which])
    bench(f"[{tag}] dgrad QKV K=2304 (B MN) +res", lambda: [ops.GemmSpec(t(Ms[s], 3 * H), wqkv, out(Ms[s], H), b_major=MN, residual=xs[s]) for s in which])
    bench(f"[{tag}] wgrad W2 [768,3072] splitK", lambda: [ops.GemmSpec(xs[s], ff[s], torch.zeros(H, FF, device=dev), a_major=MN, b_major=MN, accumulate=True, split_k=_wgrad_split(Ms[s], H, FF)) for s in which])
    bench(f"[{tag}] wgrad W1 [3072,768] splitK", lambda: [ops.GemmSpec(ff[s], xs[s], torch.zeros(FF, H, device=dev), a_major=MN, b_major=MN, accumulate=True, split_k=_wgrad_split(Ms[s], FF, H)) for s in which])
    bench(f"[{tag}] wgrad Wo [768,768] splitK", lambda: [ops.GemmSpec(xs[s], xs[s], torch.zeros(H, H, device=dev), a_major=MN, b_major=MN, accumulate=True, split_k=_wgrad_split(Ms[s], H, H)) for s in which])
    bench(f"[{tag}] wgrad Wqkv [2304,768] splitK", lambda: [ops.GemmSpec(t(Ms[s], 3 * H), xs[s], torch.zeros(3 * H, H, device=dev), a_major=MN, b_major=MN, accumulate=True, split_k=_wgrad_split(Ms[s], 3 * H, H)) for s in which])
# big square reference point
a, b = t(8192, 8192), t(8192, 8192)
bench("8192^3 plain bf16 out", lambda: [ops.GemmSpec(a, b, out(8192, 8192))], reps=5)
if ONLY:
    sys.exit(0)
c = torch.empty(8192, 8192, device=dev, dtype=bf)
torch.matmul(a, b.t(), out=c); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): torch.matmul(a, b.t(), out=c)
e1.record(); torch.cuda.synchronize()
print(f"cuBLAS 8192^3: {2*8192**3*5/(e0.elapsed_time(e1)*1e-3)/1e12:.1f} TFLOP/s")
