"""Runs attention cases one per subprocess with a timeout (a deadlocked kernel cannot take the whole session down) and
compares against a torch fp32 reference.  Usage: python tools/attn_debug.py [case ...]   case = B,Lq,Lk,masked,drop,backend"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
DEFAULT = ["2,325,32,1,0,1", "2,325,32,0,0,1", "2,32,325,0,0,1", "2,325,325,0,0,1", "64,325,325,0,0,1", "64,325,325,0,0.1,1",
           "2,130,100,1,0,1", "2,384,384,1,0,1", "1,5,7,0,0,1", "3,32,32,1,0,1", "64,325,32,1,0.1,1", "64,32,325,0,0.1,1"]

def child(case):
    import math, torch
    from m3ae_b200 import ops
    B, Lq, Lk, masked, drop, backend = case.split(",")
    B, Lq, Lk, masked, backend, drop = int(B), int(Lq), int(Lk), int(masked), int(backend), float(drop)
    H, dh, dev = 12, 64, "cuda"
    ops.set_attention_backend(backend)
    g = torch.Generator().manual_seed(0)
    qkv_q = torch.randn(B * Lq, 3 * H * dh, generator=g).to(dev).to(torch.bfloat16)
    qkv_k = torch.randn(B * Lk, 3 * H * dh, generator=g).to(dev).to(torch.bfloat16)
    q, k, v = qkv_q[:, :768], qkv_k[:, 768:1536], qkv_k[:, 1536:]
    mask = None
    if masked:
        lens = torch.randint(max(1, Lk // 4), Lk + 1, (B,), generator=g)
        mask = ((torch.arange(Lk)[None, :] >= lens[:, None]).float() * -10000.0).to(dev).contiguous()
    out = torch.empty(B * Lq, 768, dtype=torch.bfloat16, device=dev)
    lse = torch.empty(B, H, Lq, device=dev)
    kw = dict(B=B, H=H, Lq=Lq, Lk=Lk, dh=dh, mask=mask, drop_p=drop, drop_seed=5, drop_stream=1)
    dm = None
    if drop > 0:
        dm = torch.zeros(B * H * Lq * ops.drop_mask_ld(Lk), dtype=torch.uint8, device=dev)
        kw["drop_mask"] = dm
    print("  fwd launch", flush=True)
    ops.attn_fwd(q, k, v, out, lse, **kw)
    torch.cuda.synchronize()
    print("  fwd done", flush=True)
    qr, kr, vr = (t.float().detach().clone().requires_grad_(True) for t in (q, k, v))
    qf = qr.view(B, Lq, H, dh).transpose(1, 2); kf = kr.view(B, Lk, H, dh).transpose(1, 2); vf = vr.view(B, Lk, H, dh).transpose(1, 2)
    s = qf @ kf.transpose(-1, -2) / math.sqrt(dh)
    if mask is not None: s = s + mask.view(B, 1, 1, Lk)
    p = torch.softmax(s, -1)
    if drop > 0:
        bits = dm.view(B, H, Lq, -1)
        keep = ((bits[..., None] >> torch.arange(8, device=dev, dtype=torch.uint8)) & 1).flatten(-2)[..., :Lk].float()
        print(f"  keep rate {keep.mean().item():.4f}")
        p = p * keep / (1 - drop)
    o = (p @ vf).transpose(1, 2).reshape(B * Lq, 768)
    rel = lambda a, b: ((a.float() - b.float()).norm() / b.float().norm()).item()
    print(f"  fwd rel err {rel(out, o):.3e}   lse err {rel(lse, torch.logsumexp(s, -1)):.3e}", flush=True)
    dout = torch.randn(B * Lq, 768, generator=g).to(dev).to(torch.bfloat16)
    o.backward(dout.float())
    dq_b = torch.zeros_like(qkv_q); dk_b = torch.zeros_like(qkv_k)
    delta = torch.empty(B, H, Lq, device=dev)
    print("  bwd launch", flush=True)
    ops.attn_bwd(q, k, v, out, lse, dout=dout, dq=dq_b[:, :768], dk=dk_b[:, 768:1536], dv=dk_b[:, 1536:], delta=delta, **kw)
    torch.cuda.synchronize()
    print(f"  bwd rel err dq {rel(dq_b[:, :768], qr.grad):.3e} dk {rel(dk_b[:, 768:1536], kr.grad):.3e} dv {rel(dk_b[:, 1536:], vr.grad):.3e}", flush=True)

if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--child":
        child(sys.argv[2])
    else:
        for case in (sys.argv[1:] or DEFAULT):
            print("case", case, flush=True)
            try:
                r = subprocess.run([sys.executable, __file__, "--child", case], timeout=int(os.environ.get("CASE_TIMEOUT", "60")),
                                   capture_output=True, text=True)
                print(r.stdout, end="")
                if r.returncode: print("  EXIT", r.returncode, r.stderr[-600:])
            except subprocess.TimeoutExpired as e:
                print((e.stdout or b"").decode() if isinstance(e.stdout, bytes) else (e.stdout or ""), "  TIMEOUT", flush=True)
