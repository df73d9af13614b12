"""One CLIP ResidualAttentionBlock (reference m3ae/modules/vision_encoders/clip_model.py:41-63: x + attn(ln_1(x)), then
x + mlp(ln_2(x)) with a packed in-proj MHA and a QuickGELU MLP) as ONE autograd node with a persistent workspace and cached
C-ABI launch descriptors — the pre-LN sibling of functional._CrossLayersFn.

Why: built from the composable nodes (functional.layernorm / linear / attention_packed / ffn) a block costs six autograd nodes,
four separate weight casts, ~25 tensor allocations and as many ctypes struct constructions per step; with 11 blocks the
whole-model step and the small-batch classifier forward are bound by that host time, not by the GPU (tools/whole_step_probe.py).
Here a step is one ctypes call per kernel launch.  The kernels, their arguments and their order are the composable path's, so
the activations are bit-identical to it (tests/test_dropin_gpu.py compares the two; M3AE_VIT_FUSED=0 selects the old path).

Forward (7 launches)                              kept for backward
  h1  = LN1(x)                                    h1, mean1, rstd1 (x is the caller's tensor)
  qkv = h1 Wqkv^T + b            [M, 2304]        qkv
  ctx = flash_attn(qkv)          no mask          ctx, lse
  x1  = ctx Wo^T + b + x                          x1
  h2  = LN2(x1)                                   h2, mean2, rstd2
  g   = quickgelu(u), u = h2 W1^T + b  [M, 3072]  u, g      (one GEMM, two outputs)
  out = g W2^T + b + x1
Backward: the mirror image — dgrad + wgrad of every projection grouped in one persistent GEMM launch (QuickGELU' in the
dgrad epilogue), bias gradients by column sums, both LayerNorm backwards; all parameter gradients of the block are views of ONE
zero-filled fp32 buffer.  The gradient temporaries are shared by every block of the same shape (blocks run their backward
one after the other on one stream)."""
import torch

from . import functional as Fn
from . import ops

(G1, B1, WQKV, BQKV, WO, BO, G2, B2, W1, BFC, W2, BPROJ) = range(12)
_scratch = {}  # (M, D, FF, dtype, device) -> functional._Stream holding the backward temporaries shared by all blocks


def block_params(blk):
    """The block's parameters in the order of the indices above (reference names: ln_1, attn.in_proj_*, attn.out_proj,
    ln_2, mlp.c_fc, mlp.c_proj)."""
    return [blk.ln_1.weight, blk.ln_1.bias, blk.attn.in_proj_weight, blk.attn.in_proj_bias, blk.attn.out_proj.weight,
            blk.attn.out_proj.bias, blk.ln_2.weight, blk.ln_2.bias, blk.mlp.c_fc.weight, blk.mlp.c_fc.bias,
            blk.mlp.c_proj.weight, blk.mlp.c_proj.bias]


class VitBlockState:
    """Compute-dtype copies of the block's four weight matrices (refreshed in place by ONE grouped cast; same policy as
    functional.LayerOperands: always before a forward that will be differentiated — fused optimizers do not bump version
    counters — and once more on the first forward after one) plus the block's workspace pool."""

    def __init__(self):
        self.key = None
        self.alloc_key = None
        self.cast_key = None
        self.volatile = True
        self.workspaces = []

    def operands(self, params, dtype, will_train):
        key = (dtype,) + tuple((p.data_ptr(), p._version) for p in params)
        if key == self.key and not self.volatile and not will_train:
            return self
        dev = params[WQKV].device
        D, FF = params[WQKV].shape[1], params[W1].shape[0]
        alloc_key = (dtype, dev, D, FF)
        with torch.no_grad():
            if alloc_key != self.alloc_key:
                self.wqkv = torch.empty(3 * D, D, dtype=dtype, device=dev)
                self.wo = torch.empty(D, D, dtype=dtype, device=dev)
                self.w1 = torch.empty(FF, D, dtype=dtype, device=dev)
                self.w2 = torch.empty(D, FF, dtype=dtype, device=dev)
                self.alloc_key = alloc_key
                self.cast_key = None
                self.workspaces = []  # descriptors built on the old buffers are stale
            cast_key = tuple(p.data_ptr() for p in params)
            if cast_key != self.cast_key:
                for p in params:
                    if not p.is_contiguous() or p.dtype != torch.float32:
                        raise ValueError("ViT block (B200) expects contiguous fp32 master parameters")
                self.casts = ops.CastGroup([(params[WQKV].detach(), self.wqkv), (params[WO].detach(), self.wo),
                                            (params[W1].detach(), self.w1), (params[W2].detach(), self.w2)])
                self.cast_key = cast_key
            self.casts.run()
        self.key = key
        self.volatile = will_train
        return self

    def __deepcopy__(self, memo):  # copy.deepcopy(model): the copy builds its own cache (ctypes descriptors do not copy)
        return VitBlockState()

    def __reduce__(self):
        return (VitBlockState, ())

    def acquire(self, key):
        for ws in self.workspaces:
            if not ws.busy and ws.key == key:
                ws.busy = True
                return ws
        ws = Fn._Stream(key)
        ws.busy = True
        self.workspaces = [w for w in self.workspaces if w.key == key][-3:] + [ws]
        return ws


def _forward(state, nh, eps1, eps2, B, L, x, params, will_train):
    """x: [B*L, D] contiguous in the compute dtype.  Returns the workspace; st.out is the block's output (a fresh tensor)."""
    dt = x.dtype
    dev = x.device
    M, D = x.shape
    FF = params[W1].shape[0]
    w = state.operands(params, dt, will_train)
    st = state.acquire((B, L, D, FF, dt, dev, tuple(p.data_ptr() for p in params)))
    st.M, st.D, st.FF, st.B, st.L, st.nh, st.dev, st.dt = M, D, FF, B, L, nh, dev, dt
    st.p, st.w, st.x, st.eps = params, w, x, (eps1, eps2)
    f32 = torch.float32
    if "fwd" not in st.launch:
        st.h1 = st.buf("h1", (M, D), dt, dev)
        st.qkv = st.buf("qkv", (M, 3 * D), dt, dev)
        st.ctx = st.buf("ctx", (M, D), dt, dev)
        st.lse = st.buf("lse", (B, nh, L), f32, dev)
        st.x1 = st.buf("x1", (M, D), dt, dev)
        st.h2 = st.buf("h2", (M, D), dt, dev)
        st.u = st.buf("u", (M, FF), dt, dev)
        st.g = st.buf("g", (M, FF), dt, dev)
        for n in ("mean1", "rstd1", "mean2", "rstd2"):
            setattr(st, n, st.buf(n, (M,), f32, dev))
        F_ = {}
        F_["qkv"] = Fn._Gemm([ops.GemmSpec(st.h1, w.wqkv, st.qkv, bias=params[BQKV].detach())])
        F_["o"] = Fn._Gemm([ops.GemmSpec(st.ctx, w.wo, st.x1, bias=params[BO].detach(), residual=st.x1)])  # residual: patched per call (x)
        F_["ffn1"] = Fn._Gemm([ops.GemmSpec(st.h2, w.w1, st.g, bias=params[BFC].detach(), act=ops.ACT_QUICKGELU, aux_out=st.u)])
        F_["ffn2"] = Fn._Gemm([ops.GemmSpec(st.g, w.w2, st.x1, bias=params[BPROJ].detach(), residual=st.x1)])  # out: patched per call
        F_["att"] = Fn._Attn(ops.attn_params(st.qkv[:, :D], st.qkv[:, D:2 * D], st.qkv[:, 2 * D:], st.ctx, st.lse, B=B, H=nh,
                                             Lq=L, Lk=L, dh=D // nh), False)
        st.launch["fwd"] = F_
    F_ = st.launch["fwd"]
    lib = ops.L.load()
    stream = ops._stream()
    code = ops._code(dt)
    be = ops.GEMM_AUTO
    st.out = torch.empty(M, D, dtype=dt, device=dev)  # handed to the caller: always a fresh tensor
    F_["o"].arr[0].epi.residual = x.data_ptr()
    F_["ffn2"].arr[0].epi.out = st.out.data_ptr()

    def ln_fwd(src, g, b, y, mean, rstd, eps):
        ops.L.check(lib.m3ae_layernorm_fwd(src.data_ptr(), g.data_ptr(), b.data_ptr(), y.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                                           M, D, eps, code, stream), "m3ae_layernorm_fwd")

    ln_fwd(x, params[G1], params[B1], st.h1, st.mean1, st.rstd1, eps1)      # 1. ln_1
    F_["qkv"].run(be, stream)                                               # 2. packed in-proj
    F_["att"].run(stream)                                                   # 3. attention (no mask, no dropout)
    F_["o"].run(be, stream)                                                 # 4. out-proj + residual
    ln_fwd(st.x1, params[G2], params[B2], st.h2, st.mean2, st.rstd2, eps2)  # 5. ln_2
    F_["ffn1"].run(be, stream)                                              # 6. c_fc + QuickGELU (pre-activation kept)
    F_["ffn2"].run(be, stream)                                              # 7. c_proj + residual
    return st


GRAD_LAYOUT = (("wqkv", WQKV), ("bqkv", BQKV), ("wo", WO), ("bo", BO), ("g1", G1), ("b1", B1), ("g2", G2), ("b2", B2),
               ("w1", W1), ("bfc", BFC), ("w2", W2), ("bproj", BPROJ))


def _backward(st, dout):
    """dout: [M, D] contiguous in the compute dtype.  Returns (dx, [12 parameter gradients]) and frees the workspace."""
    M, D, FF, B, L, nh, dev, dt = st.M, st.D, st.FF, st.B, st.L, st.nh, st.dev, st.dt
    p, w = st.p, st.w
    f32 = torch.float32
    MN = ops.MAJOR_MN
    # every parameter gradient of the block lives in ONE zero-filled fp32 buffer (fresh per call: autograd hands views of it
    # to the optimizer as .grad)
    flat = torch.zeros(sum(p[i].numel() for _, i in GRAD_LAYOUT), dtype=f32, device=dev)
    gb, off = {}, 0
    for name, i in GRAD_LAYOUT:
        gb[name] = flat[off:off + p[i].numel()].view(p[i].shape)
        off += p[i].numel()
    if "bwd" not in st.launch:
        skey = (M, D, FF, dt, dev)
        sc = _scratch.get(skey)
        if sc is None:
            if len(_scratch) >= 2:  # shapes change rarely (train / eval batch): keep the pool small
                _scratch.pop(next(iter(_scratch)))
            sc = _scratch[skey] = Fn._Stream(skey)
        st.scratch = sc
        du = sc.buf("du", (M, FF), dt, dev)
        dh = sc.buf("dh", (M, D), dt, dev)      # dgrad output in front of a LayerNorm backward (dh2, then dh1)
        t = sc.buf("t", (M, D), dt, dev)        # LayerNorm backward output, before the residual branch is added
        dx1 = sc.buf("dx1", (M, D), dt, dev)
        dctx = sc.buf("dctx", (M, D), dt, dev)
        dqkv = sc.buf("dqkv", (M, 3 * D), dt, dev)
        delta = sc.buf("delta", (B, nh, L), f32, dev)

        def wgrad(dy, x, dw):
            return ops.GemmSpec(dy, x, dw, a_major=MN, b_major=MN, accumulate=True, split_k=Fn._wgrad_split(dy.shape[0], dw.shape[0], dw.shape[1]))

        B_ = {}
        # dout is the caller's tensor: dx1 stands in for it while the descriptors are built (patched per call)
        B_["ffn2"] = Fn._Gemm([ops.GemmSpec(dx1, w.w2, du, b_major=MN, aux_in=st.u, dact=ops.ACT_QUICKGELU), wgrad(dx1, st.g, gb["w2"])])
        B_["ffn1"] = Fn._Gemm([ops.GemmSpec(du, w.w1, dh, b_major=MN), wgrad(du, st.h2, gb["w1"])])
        B_["o"] = Fn._Gemm([ops.GemmSpec(dx1, w.wo, dctx, b_major=MN), wgrad(dx1, st.ctx, gb["wo"])])
        B_["qkv"] = Fn._Gemm([ops.GemmSpec(dqkv, w.wqkv, dh, b_major=MN), wgrad(dqkv, st.h1, gb["wqkv"])])
        B_["att"] = Fn._Attn(ops.attn_params(st.qkv[:, :D], st.qkv[:, D:2 * D], st.qkv[:, 2 * D:], st.ctx, st.lse, B=B, H=nh, Lq=L, Lk=L,
                                             dh=D // nh, dout=dctx, dq=dqkv[:, :D], dk=dqkv[:, D:2 * D], dv=dqkv[:, 2 * D:], delta=delta), True)
        for g in B_.values():  # the descriptors own the addresses; do not pin this call's gradient buffer
            if isinstance(g, Fn._Gemm):
                g.specs = None
        st.launch["bwd"] = B_
        st.bt = (du, dh, t, dx1, dctx, dqkv)
    B_ = st.launch["bwd"]
    du, dh, t, dx1, dctx, dqkv = st.bt
    B_["ffn2"].arr[0].a = B_["ffn2"].arr[1].a = dout.data_ptr()
    B_["ffn2"].arr[1].epi.out = gb["w2"].data_ptr()
    B_["ffn1"].arr[1].epi.out = gb["w1"].data_ptr()
    B_["o"].arr[1].epi.out = gb["wo"].data_ptr()
    B_["qkv"].arr[1].epi.out = gb["wqkv"].data_ptr()
    lib = ops.L.load()
    stream = ops._stream()
    code = ops._code(dt)
    be = ops.GEMM_AUTO

    def colsum(x, out):
        ops.L.check(lib.m3ae_colsum(x.data_ptr(), out.data_ptr(), x.shape[0], x.shape[1], x.stride(0), code, stream), "m3ae_colsum")

    def ln_bwd(d, x, mean, rstd, gamma, dx, dg, db):
        ops.L.check(lib.m3ae_layernorm_bwd(d.data_ptr(), x.data_ptr(), mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr(), dx.data_ptr(),
                                           dg.data_ptr(), db.data_ptr(), None, M, D, code, None, 0.0, 0, 0, None, stream), "m3ae_layernorm_bwd")

    B_["ffn2"].run(be, stream)                                      # du = (dout W2) * quickgelu'(u); dW2 += dout^T g
    colsum(dout, gb["bproj"])
    colsum(du, gb["bfc"])
    B_["ffn1"].run(be, stream)                                      # dh2 = du W1; dW1 += du^T h2
    ln_bwd(dh, st.x1, st.mean2, st.rstd2, p[G2], t, gb["g2"], gb["b2"])
    torch.add(t, dout, out=dx1)                                     # gradient of x1: LayerNorm branch + residual branch
    B_["o"].run(be, stream)                                         # dctx = dx1 Wo; dWo += dx1^T ctx
    colsum(dx1, gb["bo"])
    B_["att"].run(stream)                                           # dqkv
    colsum(dqkv, gb["bqkv"])
    B_["qkv"].run(be, stream)                                       # dh1 = dqkv Wqkv; dWqkv += dqkv^T h1
    ln_bwd(dh, st.x, st.mean1, st.rstd1, p[G1], t, gb["g1"], gb["b1"])
    dx = torch.add(t, dx1)                                          # fresh: returned to autograd
    grads = [None] * 12
    for name, i in GRAD_LAYOUT:
        grads[i] = gb[name]
    st.x = st.p = None
    st.busy = False
    return dx, grads


class _VitBlockFn(torch.autograd.Function):
    """inputs: state, heads, eps of ln_1 / ln_2, B, L, needs_bwd, x [B*L, D], then the 12 parameters.  output: [B*L, D]."""

    @staticmethod
    def forward(ctx, state, nh, eps1, eps2, B, L, needs_bwd, x, *params):
        # needs_bwd comes from the caller: ctx.needs_input_grad is True for every parameter with requires_grad even under no_grad()
        st = _forward(state, nh, eps1, eps2, B, L, x, list(params), needs_bwd)
        out, st.out = st.out, None
        ctx.st = st if needs_bwd else None
        if not needs_bwd:
            st.x = st.p = None
            st.busy = False  # no backward will come: the workspace is free again
        return out

    @staticmethod
    def backward(ctx, dout):
        st = ctx.st
        if st is None:
            raise RuntimeError("ViT block (B200): backward through a forward that ran without gradient tracking")
        if dout.dtype != st.dt or not dout.is_contiguous():
            dout = ops.cast(dout.contiguous(), torch.empty(st.M, st.D, dtype=st.dt, device=st.dev))
        dx, grads = _backward(st, dout)
        ctx.st = None
        return (None,) * 7 + (dx,) + tuple(grads)


def vit_block(blk, x, B, L):
    """x + attn(ln_1(x)), then x + mlp(ln_2(x)) (clip_model.py:59-63) for rows x [B*L, D] in the compute dtype, driven from the
    block module's own Parameters.  The fused state (operand cache, workspaces) is kept on the module, outside its state dict."""
    if not x.is_cuda:
        raise RuntimeError("ViT block (B200) needs CUDA tensors: there is no CPU fallback")
    state = blk.__dict__.get("_m3ae_fused")
    if state is None:
        state = blk.__dict__["_m3ae_fused"] = VitBlockState()
    params = block_params(blk)
    if not x.is_contiguous():
        x = x.contiguous()
    needs_bwd = torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in params))
    return _VitBlockFn.apply(state, blk.n_head, float(blk.ln_1.eps), float(blk.ln_2.eps), B, L, needs_bwd, x, *params)
