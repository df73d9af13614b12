"""Per-phase clock64 timeline of block (0,0) of the tcgen05 attention kernels (library built with -DM3AE_ATTN_TRACE)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from m3ae_b200 import ops, _lib
lib = _lib.load()
B, H, dh = 64, 12, 64
Lq, Lk = int(sys.argv[1]), int(sys.argv[2]); drop = float(sys.argv[3]) if len(sys.argv) > 3 else 0.0
dev = "cuda"
q = torch.randn(B * Lq, 2304, device=dev).bfloat16(); kv = torch.randn(B * Lk, 2304, device=dev).bfloat16()
Q, K, V = q[:, :768], kv[:, 768:1536], kv[:, 1536:]
out = torch.empty(B * Lq, 768, device=dev, dtype=torch.bfloat16); lse = torch.empty(B, H, Lq, device=dev)
dout = torch.randn_like(out); dq = torch.empty_like(q); dkv = torch.empty_like(kv); delta = torch.empty(B, H, Lq, device=dev)
dm = torch.zeros(B * H * Lq * ops.drop_mask_ld(Lk), dtype=torch.uint8, device=dev) if drop > 0 else None
kw = dict(B=B, H=H, Lq=Lq, Lk=Lk, dh=dh, drop_p=drop, drop_seed=1, drop_stream=0, drop_mask=dm)
def trace():
    buf = (ctypes.c_longlong * 512)()
    torch.cuda.synchronize(); lib.m3ae_attn_trace.argtypes = [ctypes.c_void_p, ctypes.c_int]
    lib.m3ae_attn_trace(buf, 512); return list(buf)
for _ in range(2): ops.attn_fwd(Q, K, V, out, lse, **kw)
t = trace(); t0 = t[0]
print("FWD control: start 0, KV+Q landed", t[1] - t0)
nq = (Lq + 127) // 128
for i in range(nq):
    print(f" tile {i}: ctl SE/issueS(t+1) {t[10+8*i]-t0 if i+1<nq else '-'}  PF {t[11+8*i]-t0}  PV issued {t[12+8*i]-t0} | EW: SF {t[40+10*i]-t0} ld {t[41+10*i]-t0} pass1 {t[42+10*i]-t0} bar {t[43+10*i]-t0} OFwait {t[44+10*i]-t0} resid+SE {t[45+10*i]-t0} held {t[46+10*i]-t0} PF+readout {t[47+10*i]-t0} barB+lse {t[48+10*i]-t0}")
print(" final: OF", t[98] - t0, "readout", t[99] - t0)
for _ in range(2): ops.attn_bwd(Q, K, V, out, lse, dout=dout, dq=dq[:, :768], dk=dkv[:, 768:1536], dv=dkv[:, 1536:], delta=delta, **kw)
t = trace()
nk = ((Lk + 15) // 16 * 16 + 63) // 64
t0 = t[200]
print("BWD (relative to first SF seen by EW): kernel entry", t[100] - t0, "prologue done", t[101] - t0, "all warps done", t[102] - t0)
for it in range(nk * nq):
    print(f" it {it}: ctl sdp(it+1) {t[110+4*it]-t0 if it+1<nk*nq else '-'} PF {t[111+4*it]-t0} mma {t[112+4*it]-t0} | EW: SF {t[200+6*it]-t0} ld+SE {t[201+6*it]-t0} MDwait {t[202+6*it]-t0} compute {t[203+6*it]-t0} PF {t[204+6*it]-t0} dkv {t[205+6*it]-t0 if (it % nq == 0 and it > 0) else '-'}")
