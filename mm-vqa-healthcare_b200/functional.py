"""Forward/backward of BertCrossLayer (reference bert_model.py:445-503) as ONE autograd node per layer — or per
text/image layer PAIR — built from the C-ABI kernels in `ops`.

A "stream" is one BertCrossLayer application: hidden [B,Lq,768] attends to itself, then to `encoder` [B,Lk,768], then
the FFN.  The text and image layers of one fusion index both read the pre-update x, y (m3ae_module.py:276-277), so
they are independent: `cross_layers_apply` takes 1 or 2 streams and issues every GEMM stage as one grouped persistent
launch covering both streams (the tiny text GEMMs, M = 32·B, ride in the tail wave of the image GEMMs, M = 325·B).

Kernel schedule per stream (forward):            saved for backward
  qkv  = x·Wqkv^T + b                  [M,2304]   qkv
  ctx  = flash_attn(qkv, self mask)    [M,768]    ctx, lse1
  h1   = drop(ctx·Wo^T + b) + x                   h1, mean1, rstd1
  a1   = LN(h1)                                   a1
  qc   = a1·Wq2^T + b ; kvc = y·Wkv2^T + b        qc, kvc
  ctx2 = flash_attn(qc, kvc, cross mask)          ctx2, lse2
  h2   = drop(ctx2·Wo2^T + b) + a1 ; a2 = LN(h2)  h2, mean2, rstd2, a2
  u    = a2·W1^T + b ; g = gelu(u)     [M,3072]   gelu'(u) (stored in the `u` buffer), g   (one GEMM, two outputs)
  h3   = drop(g·W2^T + b) + a2 ; out = LN(h3)     h3, mean3, rstd3
Backward runs the mirror image: LN-bwd (+bias-grad column sums) -> dgrad GEMMs (B operand MN-major, residual and
GELU' fused in the epilogue) and wgrad GEMMs (both operands MN-major, split-K atomics into fp32) grouped per stage.
"""
import itertools

import torch

from . import ops

_seed_counter = itertools.count(1)

PARAM_NAMES = [
    "attention.self.query.weight", "attention.self.query.bias",
    "attention.self.key.weight", "attention.self.key.bias",
    "attention.self.value.weight", "attention.self.value.bias",
    "attention.output.dense.weight", "attention.output.dense.bias",
    "attention.output.LayerNorm.weight", "attention.output.LayerNorm.bias",
    "crossattention.self.query.weight", "crossattention.self.query.bias",
    "crossattention.self.key.weight", "crossattention.self.key.bias",
    "crossattention.self.value.weight", "crossattention.self.value.bias",
    "crossattention.output.dense.weight", "crossattention.output.dense.bias",
    "crossattention.output.LayerNorm.weight", "crossattention.output.LayerNorm.bias",
    "intermediate.dense.weight", "intermediate.dense.bias",
    "output.dense.weight", "output.dense.bias",
    "output.LayerNorm.weight", "output.LayerNorm.bias",
]
NP = len(PARAM_NAMES)  # 26
(WQ, BQ, WK, BK, WV, BV, WO, BO, G1, B1, WQ2, BQ2, WK2, BK2, WV2, BV2, WO2, BO2, G2, B2, W1, BI, W2, BO3, G3, B3) = range(NP)


class LayerOperands:
    """Compute-dtype operand cache of one layer's fp32 master parameters: fused [Wq;Wk;Wv] and [Wk2;Wv2] weights
    (one N=2304 / N=1536 GEMM instead of three / two) cast to bf16 (product path) or copied as fp32 (parity path).
    Buffers are allocated once and refreshed IN PLACE by one grouped cast launch, so launch descriptors that point at
    them stay valid.  Refresh policy: always before a forward that will be differentiated (the optimizer moves the
    masters between two such forwards, and fused optimizers — torch.optim.AdamW(fused=True) — do NOT bump the
    parameters' version counters, so staleness cannot be detected), and once more on the first forward after one;
    consecutive no-grad forwards re-use the cache while (address, version) of every parameter is unchanged.
    Also owns the layer's workspace pool."""

    def __init__(self):
        self.key = None
        self.alloc_key = None
        self.cast_key = None
        self.volatile = True
        self.workspaces = []

    def get(self, params, dtype, will_train=True):
        key = (dtype,) + tuple((p.data_ptr(), p._version) for p in params)
        if key == self.key and not self.volatile and not will_train:
            return self
        dev = params[0].device
        H = params[WQ].shape[1]
        FF = params[W1].shape[0]
        alloc_key = (dtype, dev, H, FF)
        with torch.no_grad():
            if alloc_key != self.alloc_key:
                def buf(rows, cols):
                    return torch.empty(rows, cols, dtype=dtype, device=dev)
                self.wqkv, self.wkv2 = buf(3 * H, H), buf(2 * H, H)
                self.bqkv = torch.empty(3 * H, dtype=torch.float32, device=dev)
                self.bkv2 = torch.empty(2 * H, dtype=torch.float32, device=dev)
                self.wo, self.wq2, self.wo2, self.w1, self.w2 = buf(H, H), buf(H, H), buf(H, H), buf(FF, H), buf(H, FF)
                self.alloc_key = alloc_key
                self.cast_key = None
                self.workspaces = []  # descriptors built on the old buffers are stale
            cast_key = tuple(p.data_ptr() for p in params)
            if cast_key != self.cast_key:
                pairs = []
                for i, idx in enumerate((WQ, WK, WV)):
                    pairs.append((params[idx].detach(), self.wqkv[i * H:(i + 1) * H]))
                    pairs.append((params[(BQ, BK, BV)[i]].detach(), self.bqkv[i * H:(i + 1) * H]))
                for i, idx in enumerate((WK2, WV2)):
                    pairs.append((params[idx].detach(), self.wkv2[i * H:(i + 1) * H]))
                    pairs.append((params[(BK2, BV2)[i]].detach(), self.bkv2[i * H:(i + 1) * H]))
                pairs += [(params[WO].detach(), self.wo), (params[WQ2].detach(), self.wq2), (params[WO2].detach(), self.wo2),
                          (params[W1].detach(), self.w1), (params[W2].detach(), self.w2)]
                for src, _ in pairs:
                    if not src.is_contiguous() or src.dtype != torch.float32:
                        raise ValueError("BertCrossLayer (B200) expects contiguous fp32 master parameters")
                self.casts = ops.CastGroup(pairs)
                self.cast_key = cast_key
            self.casts.run()
        self.key = key
        self.volatile = will_train
        return self

    def acquire(self, key):
        """A free workspace for this (shape, dtype, parameter-address) key, or a new one.  A workspace is busy from
        forward until its backward has run (or until forward returns when autograd is off)."""
        for ws in self.workspaces:
            if not ws.busy and ws.key == key:
                ws.busy = True
                return ws
        ws = _Stream(key)
        ws.busy = True
        self.workspaces = [w for w in self.workspaces if w.key == key][-3:] + [ws]  # drop stale shapes, bound the pool
        return ws


class LayerConfig:
    def __init__(self, num_heads=12, eps=1e-12, hidden_dropout=0.0, attn_dropout=0.0, compute_dtype=torch.bfloat16,
                 gemm_backend=ops.GEMM_AUTO):
        self.num_heads = num_heads
        self.eps = eps
        self.hidden_dropout = hidden_dropout
        self.attn_dropout = attn_dropout
        self.compute_dtype = compute_dtype
        self.gemm_backend = gemm_backend


def _wgrad_split(tokens, n_out, n_in):
    """split-K factor for dW[n_out,n_in] += dY^T X over `tokens`: aim at ~1 wave of 148 CTAs of 128x256 tiles."""
    tiles = ((n_out + 127) // 128) * ((n_in + 255) // 256)
    kblocks = (tokens + 63) // 64
    return max(1, min(kblocks, 148 // tiles if tiles <= 148 else 1))


def _as2d(t):
    return t.reshape(-1, t.shape[-1])


def _mask2d(mask, B, L):
    if mask is None:
        return None
    if mask.numel() != B * L:
        raise ValueError(f"attention mask must broadcast as [B,1,1,L]=[{B},1,1,{L}], got {tuple(mask.shape)}")
    m = mask.reshape(B, L)
    if m.dtype != torch.float32:
        m = m.float()
    return m.contiguous()


class _Stream:
    """Persistent workspace of one BertCrossLayer application ("stream"): the activations kept for backward, the
    gradient temporaries and the prebuilt C-ABI launch descriptors (ctypes structs) that point at them.  Re-used
    across steps so that a step costs one ctypes call per kernel launch instead of dozens of tensor allocations and
    struct constructions (host issue time was 17 ms/step before this, vs 22 ms of GPU time)."""

    def __init__(self, key):
        self.key = key
        self.busy = False
        self.t = {}
        self.launch = {}

    def buf(self, name, shape, dtype, device):
        t = self.t.get(name)
        if t is None:
            t = torch.empty(shape, dtype=dtype, device=device)
            self.t[name] = t
        return t


def _ptr(t):
    return t.data_ptr() if t is not None else None


class LazyAttentionProbs:
    """Attention probabilities [B,nh,Lq,Lk] (fp32) of one attention block, computed on first use.

    The reference's fusion loop asks every layer for its probabilities (output_attentions=True is hard-wired at
    m3ae_module.py:276-277) and then drops them unless the caller of infer() wanted them (:282-283).  Materialising them
    eagerly costs two extra kernels and a [B,12,Lq,Lk] fp32 tensor per block (324 MB per image layer at batch 64), so
    the layer returns this handle instead: it remembers the workspace that still holds q, k, the log-sum-exp and the
    dropout seed, and runs m3ae_attn_probs when something reads it (any torch function, attribute or index).  The
    workspace belongs to the layer and is re-used by the layer's NEXT forward; reading the handle after that raises."""

    def __init__(self, ws, which, gen):
        self._ws, self._which, self._gen = ws, which, gen
        self._value = None

    def materialize(self):
        if self._value is None:
            st = self._ws
            if st.gen != self._gen:
                raise RuntimeError("attention probabilities were requested after the layer ran again: read (or .materialize()) "
                                   "them before the next forward of the same layer")
            nh = st.nh
            H = st.H
            if self._which == 0:
                q, k, lse, Lk, mask, site = st.qkv[:, :H], st.qkv[:, H:2 * H], st.lse1, st.Lq, st.mask, 0
            else:
                q, k, lse, Lk, mask, site = st.qc, st.kvc[:, :H], st.lse2, st.Lk, st.enc_mask, 2
            probs = torch.empty(st.B, nh, st.Lq, Lk, dtype=torch.float32, device=st.dev)
            with torch.no_grad():
                ops.attn_probs(q, k, lse, probs, B=st.B, H=nh, Lq=st.Lq, Lk=Lk, dh=st.dh, mask=mask, drop_p=st.pa,
                               drop_seed=st.seed, drop_stream=site)
            self._value = probs
            self._ws = None
        return self._value

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        def conv(a):
            if isinstance(a, LazyAttentionProbs):
                return a.materialize()
            if isinstance(a, (list, tuple)):
                return type(a)(conv(x) for x in a)
            return a
        return func(*conv(args), **{k: conv(v) for k, v in (kwargs or {}).items()})

    def __getattr__(self, name):  # only reached for names this class does not define: delegate to the tensor
        if name.startswith("_"):
            raise AttributeError(name)
        return getattr(self.materialize(), name)

    def __getitem__(self, idx):
        return self.materialize()[idx]

    def __len__(self):
        return len(self.materialize())

    def __repr__(self):
        return f"LazyAttentionProbs(materialized={self._value is not None})"


class _Gemm:
    """A cached grouped-GEMM launch: the ctypes problem array is built once; `patch` lists (problem index, field path,
    getter) for operands whose address is not owned by the workspace (layer inputs, incoming gradients)."""

    def __init__(self, specs):
        self.specs = specs  # keeps the tensors alive
        self.n = len(specs)
        self.arr = (ops.L.GemmProblem * self.n)(*[s.problem for s in specs])
        self.code = ops._code(specs[0].in_dtype)
        self.fn = ops.L.load().m3ae_gemm

    def run(self, be, stream):
        ops.L.check(self.fn(self.arr, self.n, self.code, be, stream), "m3ae_gemm")


class _Attn:
    def __init__(self, params, bwd):
        self.p = params
        lib = ops.L.load()
        self.fn = lib.m3ae_attn_bwd if bwd else lib.m3ae_attn_fwd
        self.ref = ops.C.byref(params)

    def run(self, stream):
        ops.L.check(self.fn(self.ref, stream), "m3ae_attn")


def _forward_streams(cfg, training, streams, will_train=True):
    """streams: list of dicts(hidden, encoder, mask, enc_mask, params(list of 26), operands(LayerOperands))."""
    dt = cfg.compute_dtype
    be = cfg.gemm_backend
    nh = cfg.num_heads
    pa = cfg.attn_dropout if training else 0.0
    ph = cfg.hidden_dropout if training else 0.0
    stream = ops._stream()
    S = []
    for s in streams:
        h, e = s["hidden"], s["encoder"]
        B, Lq, H = h.shape
        Lk = e.shape[1]
        params = s["params"]
        dev = h.device
        key = (B, Lq, Lk, H, dt, dev, pa, ph, s["mask"] is None, s["enc_mask"] is None,
               tuple(p.data_ptr() for p in params))
        w = s["operands"].get(params, dt, will_train)
        st = s["operands"].acquire(key)
        st.B, st.Lq, st.H, st.Lk = B, Lq, H, Lk
        st.gen = getattr(st, "gen", 0) + 1  # LazyAttentionProbs handles of earlier forwards become invalid
        st.nh = nh
        st.dh = H // nh
        st.M, st.Mk = B * Lq, B * Lk
        st.dev = dev
        st.p, st.w = params, w
        st.FF = params[W1].shape[0]
        st.x = _as2d(h) if h.dtype == dt and h.is_contiguous() else ops.cast(h.contiguous(), st.buf("x_cast", (st.M, H), dt, dev))
        st.y = _as2d(e) if e.dtype == dt and e.is_contiguous() else ops.cast(e.contiguous(), st.buf("y_cast", (st.Mk, H), dt, dev))
        st.mask = _mask2d(s["mask"], B, Lq)
        st.enc_mask = _mask2d(s["enc_mask"], B, Lk)
        st.seed = (torch.initial_seed() + 0x9E3779B97F4A7C15 * next(_seed_counter)) & 0xFFFFFFFFFFFFFFFF if training else 0
        st.pa, st.ph = pa, ph
        S.append(st)
    S0 = S[0]
    gid = tuple(id(st) for st in S)  # descriptors of a grouped launch live on the first stream, keyed by its partners

    if S0.launch.get("fwd_gid") != gid:
        for st in S:
            st.launch = {}
        S0.launch["fwd_gid"] = gid
        f32 = torch.float32
        for st in S:
            H, M, Mk, FF, dev = st.H, st.M, st.Mk, st.FF, st.dev
            st.qkv = st.buf("qkv", (M, 3 * H), dt, dev)
            st.ctx = st.buf("ctx", (M, H), dt, dev)
            st.lse1 = st.buf("lse1", (st.B, nh, st.Lq), f32, dev)
            st.dm1 = st.dm2 = None
            if pa > 0.0 and dt == torch.bfloat16:  # dropout keep-bit caches: written by forward, read by backward
                st.dm1 = st.buf("dm1", (st.B * nh * st.Lq * ops.drop_mask_ld(st.Lq),), torch.uint8, dev)
                st.dm2 = st.buf("dm2", (st.B * nh * st.Lq * ops.drop_mask_ld(st.Lk),), torch.uint8, dev)
            # keep bits of the three hidden dropouts: written by the GEMM epilogues, read by the LayerNorm backward
            st.hb1 = st.hb2 = st.hb3 = None
            if ph > 0.0 and dt == torch.bfloat16 and be != ops.GEMM_SIMT and H % 32 == 0:
                st.hb1, st.hb2, st.hb3 = (st.buf(n, (M, H // 32), torch.int32, dev) for n in ("hb1", "hb2", "hb3"))
            st.h1 = st.buf("h1", (M, H), dt, dev)
            st.a1 = st.buf("a1", (M, H), dt, dev)
            st.mean1, st.rstd1 = st.buf("mean1", (M,), f32, dev), st.buf("rstd1", (M,), f32, dev)
            st.qc = st.buf("qc", (M, H), dt, dev)
            st.kvc = st.buf("kvc", (Mk, 2 * H), dt, dev)
            st.ctx2 = st.buf("ctx2", (M, H), dt, dev)
            st.lse2 = st.buf("lse2", (st.B, nh, st.Lq), f32, dev)
            st.h2 = st.buf("h2", (M, H), dt, dev)
            st.a2 = st.buf("a2", (M, H), dt, dev)
            st.mean2, st.rstd2 = st.buf("mean2", (M,), f32, dev), st.buf("rstd2", (M,), f32, dev)
            st.u = st.buf("u", (M, FF), dt, dev)
            st.g = st.buf("g", (M, FF), dt, dev)
            st.h3 = st.buf("h3", (M, H), dt, dev)
            st.mean3, st.rstd3 = st.buf("mean3", (M,), f32, dev), st.buf("rstd3", (M,), f32, dev)
        L_ = S0.launch
        L_["qkv"] = _Gemm([ops.GemmSpec(st.x, st.w.wqkv, st.qkv, bias=st.w.bqkv) for st in S])
        L_["o1"] = _Gemm([ops.GemmSpec(st.ctx, st.w.wo, st.h1, bias=st.p[BO], residual=st.x, drop_p=ph, drop_stream=1, drop_bits=st.hb1) for st in S])
        specs = []
        for st in S:
            specs.append(ops.GemmSpec(st.a1, st.w.wq2, st.qc, bias=st.p[BQ2]))
            specs.append(ops.GemmSpec(st.y, st.w.wkv2, st.kvc, bias=st.w.bkv2))
        L_["xproj"] = _Gemm(specs)
        L_["o2"] = _Gemm([ops.GemmSpec(st.ctx2, st.w.wo2, st.h2, bias=st.p[BO2], residual=st.a1, drop_p=ph, drop_stream=3, drop_bits=st.hb2) for st in S])
        L_["ffn1"] = _Gemm([ops.GemmSpec(st.a2, st.w.w1, st.g, bias=st.p[BI], act=ops.ACT_GELU, aux_out=st.u, aux_is_grad=True) for st in S])
        L_["ffn2"] = _Gemm([ops.GemmSpec(st.g, st.w.w2, st.h3, bias=st.p[BO3], residual=st.a2, drop_p=ph, drop_stream=4, drop_bits=st.hb3) for st in S])
        for st in S:
            H = st.H
            st.launch["att1"] = _Attn(ops.attn_params(st.qkv[:, :H], st.qkv[:, H:2 * H], st.qkv[:, 2 * H:], st.ctx, st.lse1,
                                                      B=st.B, H=nh, Lq=st.Lq, Lk=st.Lq, dh=st.dh, mask=st.mask, drop_p=pa,
                                                      drop_stream=0, drop_mask=st.dm1), False)
            st.launch["att2"] = _Attn(ops.attn_params(st.qc, st.kvc[:, :H], st.kvc[:, H:], st.ctx2, st.lse2, B=st.B, H=nh,
                                                      Lq=st.Lq, Lk=st.Lk, dh=st.dh, mask=st.enc_mask, drop_p=pa,
                                                      drop_stream=2, drop_mask=st.dm2), False)
    L_ = S0.launch
    lib = ops.L.load()

    def ln_fwd(x, g, b, y, mean, rstd, M, H):
        ops.L.check(lib.m3ae_layernorm_fwd(x.data_ptr(), g.data_ptr(), b.data_ptr(), y.data_ptr(), mean.data_ptr(),
                                           rstd.data_ptr(), M, H, cfg.eps, ops._code(dt), stream), "m3ae_layernorm_fwd")

    # external addresses / per-call values
    for i, st in enumerate(S):
        L_["qkv"].arr[i].a = st.x.data_ptr()
        L_["o1"].arr[i].epi.residual = st.x.data_ptr()
        L_["xproj"].arr[2 * i + 1].a = st.y.data_ptr()
        for nm in ("o1", "o2", "ffn2"):
            L_[nm].arr[i].epi.drop_seed = st.seed
        a1, a2 = st.launch["att1"].p, st.launch["att2"].p
        a1.mask, a2.mask = _ptr(st.mask), _ptr(st.enc_mask)
        a1.drop_seed = a2.drop_seed = st.seed

    L_["qkv"].run(be, stream)                                  # 1. fused QKV projection
    for st in S:
        st.launch["att1"].run(stream)                          # 2. self attention
    L_["o1"].run(be, stream)                                   # 3. out-proj + dropout + residual
    for st in S:
        ln_fwd(st.h1, st.p[G1], st.p[B1], st.a1, st.mean1, st.rstd1, st.M, st.H)   # 4. LayerNorm
    L_["xproj"].run(be, stream)                                # 5. cross Q (own stream) and fused K|V (other stream)
    for st in S:
        st.launch["att2"].run(stream)                          # 6. cross attention
    L_["o2"].run(be, stream)                                   # 7. cross out-proj + dropout + residual
    for st in S:
        ln_fwd(st.h2, st.p[G2], st.p[B2], st.a2, st.mean2, st.rstd2, st.M, st.H)   # 8.
    L_["ffn1"].run(be, stream)                                 # 9. FFN up + GELU (pre-activation kept)
    L_["ffn2"].run(be, stream)                                 # 10. FFN down + dropout + residual
    for st in S:
        st.out = torch.empty(st.M, st.H, dtype=dt, device=st.dev)  # returned to the caller: always a fresh tensor
        ln_fwd(st.h3, st.p[G3], st.p[B3], st.out, st.mean3, st.rstd3, st.M, st.H)  # 11.
    return S


def _backward_streams(cfg, S, douts):
    """S: forward workspaces; douts: list of [M,H] compute-dtype grads. Returns per stream (dx, dy, [26 param grads])."""
    dt = cfg.compute_dtype
    be = cfg.gemm_backend
    nh = cfg.num_heads
    MN = ops.MAJOR_MN
    stream = ops._stream()
    S0 = S[0]
    lib = ops.L.load()
    f32 = torch.float32

    def wgrad(dy, x, dw):
        return ops.GemmSpec(dy, x, dw, a_major=MN, b_major=MN, accumulate=True,
                            split_k=_wgrad_split(dy.shape[0], dw.shape[0], dw.shape[1]))

    for st in S:
        H, FF, dev = st.H, st.FF, st.dev
        # all parameter gradients of the layer live in ONE zero-filled fp32 buffer (one memset instead of ~30 fills);
        # fused projections get contiguous blocks so their wgrad GEMM writes q|k|v (k|v) gradients in one go.  It is a
        # fresh tensor every call: autograd hands views of it to the optimizer as .grad.
        sizes = [("wqkv", 3 * H * H), ("bqkv", 3 * H), ("wo", H * H), ("bo", H), ("g1", H), ("b1", H), ("wq2", H * H),
                 ("bq2", H), ("wkv2", 2 * H * H), ("bkv2", 2 * H), ("wo2", H * H), ("bo2", H), ("g2", H), ("b2", H),
                 ("w1", FF * H), ("bi", FF), ("w2", H * FF), ("bo3", H), ("g3", H), ("b3", H)]
        flat = torch.zeros(sum(n for _, n in sizes), dtype=f32, device=dev)
        gb, off = {}, 0
        for name, n in sizes:
            gb[name] = flat[off:off + n]
            off += n
        st.gb = gb
        st.dx = torch.empty(st.M, H, dtype=dt, device=dev)   # returned to autograd: fresh
        st.dy = torch.empty(st.Mk, H, dtype=dt, device=dev)

    if "bwd" not in S0.launch:
        for st in S:
            H, FF, M, Mk, dev = st.H, st.FF, st.M, st.Mk, st.dev
            for name, shape in (("dh3", (M, H)), ("dh2", (M, H)), ("dh1", (M, H)), ("du", (M, FF)), ("da2", (M, H)),
                                ("dctx2", (M, H)), ("dqc", (M, H)), ("dkvc", (Mk, 2 * H)), ("da1", (M, H)), ("dctx", (M, H)),
                                ("dqkv", (M, 3 * H))):
                setattr(st, name, st.buf(name, shape, dt, dev))
            for name in ("dh3d", "dh2d", "dh1d"):
                setattr(st, name, st.buf(name, (M, H), dt, dev) if st.ph > 0 else getattr(st, name[:-1]))
            st.delta = st.buf("delta", (st.B, nh, st.Lq), f32, dev)
        B_ = {}
        gbv = lambda st, n, r, c: st.gb[n].view(r, c)  # noqa: E731
        B_["ffn2"] = _Gemm([ops.GemmSpec(st.dh3d, st.w.w2, st.du, b_major=MN, aux_in=st.u, dact=ops.ACT_GELU, aux_is_grad=True) for st in S] +
                           [wgrad(st.dh3d, st.g, gbv(st, "w2", st.H, st.FF)) for st in S])
        B_["ffn1"] = _Gemm([ops.GemmSpec(st.du, st.w.w1, st.da2, b_major=MN, residual=st.dh3) for st in S] +
                           [wgrad(st.du, st.a2, gbv(st, "w1", st.FF, st.H)) for st in S])
        B_["o2"] = _Gemm([ops.GemmSpec(st.dh2d, st.w.wo2, st.dctx2, b_major=MN) for st in S] +
                         [wgrad(st.dh2d, st.ctx2, gbv(st, "wo2", st.H, st.H)) for st in S])
        B_["xd"] = _Gemm([ops.GemmSpec(st.dqc, st.w.wq2, st.da1, b_major=MN, residual=st.dh2) for st in S] +
                         [ops.GemmSpec(st.dkvc, st.w.wkv2, st.dy, b_major=MN) for st in S])
        B_["xw"] = _Gemm([wgrad(st.dqc, st.a1, gbv(st, "wq2", st.H, st.H)) for st in S] +
                         [wgrad(st.dkvc, st.y, gbv(st, "wkv2", 2 * st.H, st.H)) for st in S])
        B_["o1"] = _Gemm([ops.GemmSpec(st.dh1d, st.w.wo, st.dctx, b_major=MN) for st in S] +
                         [wgrad(st.dh1d, st.ctx, gbv(st, "wo", st.H, st.H)) for st in S])
        B_["qkv"] = _Gemm([ops.GemmSpec(st.dqkv, st.w.wqkv, st.dx, b_major=MN, residual=st.dh1) for st in S] +
                          [wgrad(st.dqkv, st.x, gbv(st, "wqkv", 3 * st.H, st.H)) for st in S])
        S0.launch["bwd"] = B_
        for st in S:
            H = st.H
            st.launch["batt2"] = _Attn(ops.attn_params(
                st.qc, st.kvc[:, :H], st.kvc[:, H:], st.ctx2, st.lse2, B=st.B, H=nh, Lq=st.Lq, Lk=st.Lk, dh=st.dh,
                mask=st.enc_mask, drop_p=st.pa, drop_stream=2, dout=st.dctx2, dq=st.dqc, dk=st.dkvc[:, :H],
                dv=st.dkvc[:, H:], delta=st.delta, drop_mask=st.dm2), True)
            st.launch["batt1"] = _Attn(ops.attn_params(
                st.qkv[:, :H], st.qkv[:, H:2 * H], st.qkv[:, 2 * H:], st.ctx, st.lse1, B=st.B, H=nh, Lq=st.Lq, Lk=st.Lq,
                dh=st.dh, mask=st.mask, drop_p=st.pa, drop_stream=0, dout=st.dctx, dq=st.dqkv[:, :H],
                dk=st.dqkv[:, H:2 * H], dv=st.dqkv[:, 2 * H:], delta=st.delta, drop_mask=st.dm1), True)
    B_ = S0.launch["bwd"]
    n = len(S)

    # external addresses: this call's gradient buffers, the layer inputs, masks, seeds
    for i, st in enumerate(S):
        gb = st.gb
        B_["ffn2"].arr[n + i].epi.out = gb["w2"].data_ptr()
        B_["ffn1"].arr[n + i].epi.out = gb["w1"].data_ptr()
        B_["o2"].arr[n + i].epi.out = gb["wo2"].data_ptr()
        B_["xd"].arr[n + i].epi.out = st.dy.data_ptr()
        B_["xw"].arr[i].epi.out = gb["wq2"].data_ptr()
        B_["xw"].arr[n + i].epi.out = gb["wkv2"].data_ptr()
        B_["xw"].arr[n + i].b = st.y.data_ptr()
        B_["o1"].arr[n + i].epi.out = gb["wo"].data_ptr()
        B_["qkv"].arr[i].epi.out = st.dx.data_ptr()
        B_["qkv"].arr[n + i].epi.out = gb["wqkv"].data_ptr()
        B_["qkv"].arr[n + i].b = st.x.data_ptr()
        for nm in ("batt1", "batt2"):
            st.launch[nm].p.drop_seed = st.seed
        st.launch["batt1"].p.mask = _ptr(st.mask)
        st.launch["batt2"].p.mask = _ptr(st.enc_mask)

    code = ops._code(dt)

    def ln_bwd(st, d, h, mean, rstd, gamma, dg, db, dbias, dx, dxd, site, bits):
        """LayerNorm backward; dxd = gradient pushed back through the hidden dropout of the GEMM that produced h (same
        Philox stream as the forward epilogue); dbias accumulates that GEMM's bias gradient."""
        ops.L.check(lib.m3ae_layernorm_bwd(d.data_ptr(), h.data_ptr(), mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr(),
                                           dx.data_ptr(), dg.data_ptr(), db.data_ptr(), dbias.data_ptr(), st.M, st.H, code,
                                           dxd.data_ptr() if st.ph > 0 else None, st.ph, st.seed, site, _ptr(bits), stream),
                    "m3ae_layernorm_bwd")

    def colsum(x, out):
        ops.L.check(lib.m3ae_colsum(x.data_ptr(), out.data_ptr(), x.shape[0], x.shape[1], x.stride(0), code, stream), "m3ae_colsum")

    for st, d in zip(S, douts):   # ---- output LayerNorm + FFN
        ln_bwd(st, d, st.h3, st.mean3, st.rstd3, st.p[G3], st.gb["g3"], st.gb["b3"], st.gb["bo3"], st.dh3, st.dh3d, 4, st.hb3)
    B_["ffn2"].run(be, stream)
    for st in S:
        colsum(st.du, st.gb["bi"])
    B_["ffn1"].run(be, stream)
    for st in S:                  # ---- cross-attention block
        ln_bwd(st, st.da2, st.h2, st.mean2, st.rstd2, st.p[G2], st.gb["g2"], st.gb["b2"], st.gb["bo2"], st.dh2, st.dh2d, 3, st.hb2)
    B_["o2"].run(be, stream)
    for st in S:
        st.launch["batt2"].run(stream)
        colsum(st.dqc, st.gb["bq2"])
        colsum(st.dkvc, st.gb["bkv2"])
    B_["xd"].run(be, stream)
    B_["xw"].run(be, stream)
    for st in S:                  # ---- self-attention block
        ln_bwd(st, st.da1, st.h1, st.mean1, st.rstd1, st.p[G1], st.gb["g1"], st.gb["b1"], st.gb["bo"], st.dh1, st.dh1d, 1, st.hb1)
    B_["o1"].run(be, stream)
    for st in S:
        st.launch["batt1"].run(stream)
        colsum(st.dqkv, st.gb["bqkv"])
    B_["qkv"].run(be, stream)
    res = []
    for st in S:
        H, FF, gb = st.H, st.FF, st.gb
        g_ = [None] * NP
        wqkv, bqkv, wkv2, bkv2 = gb["wqkv"].view(3 * H, H), gb["bqkv"], gb["wkv2"].view(2 * H, H), gb["bkv2"]
        g_[WQ], g_[WK], g_[WV] = wqkv[:H], wqkv[H:2 * H], wqkv[2 * H:]
        g_[BQ], g_[BK], g_[BV] = bqkv[:H], bqkv[H:2 * H], bqkv[2 * H:]
        g_[WO], g_[BO], g_[G1], g_[B1] = gb["wo"].view(H, H), gb["bo"], gb["g1"], gb["b1"]
        g_[WQ2], g_[BQ2] = gb["wq2"].view(H, H), gb["bq2"]
        g_[WK2], g_[WV2], g_[BK2], g_[BV2] = wkv2[:H], wkv2[H:], bkv2[:H], bkv2[H:]
        g_[WO2], g_[BO2], g_[G2], g_[B2] = gb["wo2"].view(H, H), gb["bo2"], gb["g2"], gb["b2"]
        g_[W1], g_[BI], g_[W2], g_[BO3] = gb["w1"].view(FF, H), gb["bi"], gb["w2"].view(H, FF), gb["bo3"]
        g_[G3], g_[B3] = gb["g3"], gb["b3"]
        res.append((st.dx, st.dy, g_))
        st.gb = None
        st.busy = False  # workspace may be re-used by the next forward
    return res


class _CrossLayersFn(torch.autograd.Function):
    """inputs: cfg, training, holder (list that receives the forward workspaces, or None), operands(list), n_streams,
    needs_bwd, then per stream: hidden, encoder, mask, enc_mask, 26 params.  outputs: per stream out [B,Lq,H]."""

    @staticmethod
    def forward(ctx, cfg, training, holder, operands, n_streams, needs_bwd, *flat):
        per = 4 + NP
        streams = []
        for i in range(n_streams):
            c = flat[i * per:(i + 1) * per]
            streams.append(dict(hidden=c[0], encoder=c[1], mask=c[2], enc_mask=c[3], params=list(c[4:]), operands=operands[i]))
        # needs_bwd comes from the caller (cross_layers_apply): ctx.needs_input_grad is True for every parameter with
        # requires_grad even under torch.no_grad(), so it cannot tell an eval forward from a training one
        S = _forward_streams(cfg, training, streams, will_train=needs_bwd)
        ctx.cfg = cfg
        ctx.S = S if needs_bwd else None
        ctx.in_dtypes = [(s["hidden"].dtype, s["encoder"].dtype) for s in streams]
        outs = []
        for st, s in zip(S, streams):
            o = st.out.view(st.B, st.Lq, st.H)
            if o.dtype != s["hidden"].dtype:
                o = ops.cast(st.out, torch.empty(st.B, st.Lq, st.H, dtype=s["hidden"].dtype, device=o.device))
            outs.append(o)
            st.out = None
            if not needs_bwd:
                st.busy = False  # no backward will come: the workspace is free again
        if holder is not None:
            holder.extend(S)
        return tuple(outs)

    @staticmethod
    def backward(ctx, *grads):
        cfg, S = ctx.cfg, ctx.S
        if S is None:
            raise RuntimeError("BertCrossLayer (B200): backward through a forward that ran without gradient tracking")
        dt = cfg.compute_dtype
        douts = []
        for i, st in enumerate(S):
            g = grads[i]
            if g is None:
                g = torch.zeros(st.M, st.H, dtype=dt, device=st.dev)
            g = g.reshape(st.M, st.H)
            if g.dtype != dt or not g.is_contiguous():
                g = ops.cast(g.contiguous(), torch.empty(st.M, st.H, dtype=dt, device=g.device))
            douts.append(g)
        res = _backward_streams(cfg, S, douts)
        flat = [None, None, None, None, None, None]
        for (dx, dy, pg), st, (hd, ed) in zip(res, S, ctx.in_dtypes):
            dx = dx.view(st.B, st.Lq, st.H)
            dy = dy.view(st.B, st.Lk, st.H)
            if dx.dtype != hd:
                dx = ops.cast(dx, torch.empty_like(dx, dtype=hd))
            if dy.dtype != ed:
                dy = ops.cast(dy, torch.empty_like(dy, dtype=ed))
            flat += [dx, dy, None, None] + pg
        ctx.S = None
        return tuple(flat)


def cross_layers_apply(cfg, training, streams, want_probs=False):
    """streams: list (1 or 2) of dicts(hidden, encoder, mask, enc_mask, params, operands).
    Returns per stream a tuple (out,) or (out, self_probs, cross_probs) — the probabilities as LazyAttentionProbs handles
    (non-differentiable, computed when read)."""
    if not 1 <= len(streams) <= 2:
        raise ValueError("cross_layers_apply takes one layer or one text/image layer pair")
    flat = []
    for s in streams:
        if not s["hidden"].is_cuda:
            raise RuntimeError("BertCrossLayer (B200) needs CUDA tensors: the fusion path has no CPU fallback")
        flat += [s["hidden"], s["encoder"], s["mask"], s["enc_mask"]] + list(s["params"])
    needs_bwd = torch.is_grad_enabled() and any(torch.is_tensor(t) and t.requires_grad for t in flat)
    holder = [] if want_probs else None
    outs = _CrossLayersFn.apply(cfg, training, holder, [s["operands"] for s in streams], len(streams), needs_bwd, *flat)
    if not want_probs:
        return [(o,) for o in outs]
    return [(o, LazyAttentionProbs(st, 0, st.gen), LazyAttentionProbs(st, 1, st.gen)) for o, st in zip(outs, holder)]


# ----------------------------------------------------------------------------- composable autograd pieces
# The uni-modal encoders (CLIP ViT-B/16, RoBERTa-base), the generative decoder head and the small heads (poolers,
# vqa_head, losses) are built from these: each is one autograd node over the C-ABI kernels.  Epilogue order of every GEMM
# = the reference modules': (+bias) -> act -> dropout -> (+residual).
def new_seed():
    """A fresh 64-bit dropout seed (Philox key) derived from torch's seed: deterministic under torch.manual_seed."""
    return (torch.initial_seed() + 0x9E3779B97F4A7C15 * next(_seed_counter)) & 0xFFFFFFFFFFFFFFFF


def _as_compute(t, dtype, shape=None):
    """2-D contiguous view of `t` in the compute dtype (cast kernel when needed)."""
    t2 = t if shape is None else t.reshape(shape)
    if t2.dtype == dtype and t2.is_contiguous():
        return t2
    return ops.cast(t2.contiguous(), torch.empty(t2.shape, dtype=dtype, device=t2.device))


class _LinearFn(torch.autograd.Function):
    """y = drop(act(x W^T + b (+ extra_bias))) (+ residual) with W fp32 master [N,K]; compute in `dtype` (bf16 tcgen05 /
    fp32 SIMT).  Replaces nn.Linear at m3ae_module.py:235,252 (modality projections; extra_bias = modality type embedding
    row, :260-263), the decoder / CLIP / RoBERTa projections with their residual adds and hidden dropouts
    (m3ae_decoder.py:62,80,136; clip_model.py:58-62; HF RobertaSelfOutput / RobertaOutput)."""

    @staticmethod
    def forward(ctx, x, weight, bias, extra_bias, residual, act, dtype, out_dtype, drop_p, drop_seed, drop_site):
        K = x.shape[-1]
        N = weight.shape[0]
        M = x.numel() // K
        dev = x.device
        xc = _as_compute(x, dtype, (M, K))
        wc = _as_compute(weight.detach(), dtype)
        b = None
        if bias is not None:
            b = bias.detach().float().contiguous()
            if extra_bias is not None:
                b = ops.add_rowvec(b.view(1, N), extra_bias.detach().float().contiguous(), torch.empty(1, N, dtype=torch.float32, device=dev)).view(N)
        elif extra_bias is not None:
            b = extra_bias.detach().float().contiguous()
        rc = None
        if residual is not None:
            rc = residual.reshape(M, N)
            if rc.dtype != out_dtype or not rc.is_contiguous():
                rc = ops.cast(rc.contiguous(), torch.empty(M, N, dtype=out_dtype, device=dev))
        out = torch.empty(M, N, dtype=out_dtype, device=dev)
        ops.gemm(xc, wc, out, bias=b, act=act, residual=rc, drop_p=drop_p, drop_seed=drop_seed, drop_stream=drop_site)
        ctx.act = act
        ctx.save_for_backward(xc, wc)
        ctx.meta = (x.shape, x.dtype, dtype, bias is not None, extra_bias is not None, weight.dtype, residual is not None,
                    None if residual is None else (residual.shape, residual.dtype), drop_p, drop_seed, drop_site, out_dtype)
        return out.view(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        if ctx.act != ops.ACT_NONE:
            raise NotImplementedError("backward through a fused activation is provided by functional.ffn / activation()")
        xc, wc = ctx.saved_tensors
        xshape, xdtype, dtype, has_b, has_e, wdtype, has_res, res_meta, drop_p, drop_seed, drop_site, out_dtype = ctx.meta
        M, K = xc.shape
        N = wc.shape[0]
        dev = dy.device
        d2 = dy.reshape(M, N)
        if drop_p > 0.0:  # gradient pushed back through the epilogue dropout (same Philox stream, element = row*N + col)
            d2c = d2.contiguous()
            d2 = ops.dropout_apply(d2c, torch.empty_like(d2c), drop_p, drop_seed, drop_site)
        if dtype == torch.bfloat16 and N % 8 != 0:
            # TMA needs 16-byte row pitches: stage dY in a buffer whose leading dimension is padded (vocab 30522 -> 30528)
            pad = torch.zeros(M, (N + 7) // 8 * 8, dtype=dtype, device=dev)
            pad[:, :N].copy_(d2)
            dyc = pad[:, :N]
        else:
            dyc = _as_compute(d2, dtype)
        need_x, need_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        dx = dw = None
        specs = []
        if need_x:
            dxc = torch.empty(M, K, dtype=dtype, device=dev)
            specs.append(ops.GemmSpec(dyc, wc, dxc, b_major=ops.MAJOR_MN))
        if need_w:
            dw = torch.zeros(N, K, dtype=torch.float32, device=dev)
            specs.append(ops.GemmSpec(dyc, xc, dw, a_major=ops.MAJOR_MN, b_major=ops.MAJOR_MN, accumulate=True,
                                      split_k=_wgrad_split(M, N, K)))
        if specs:
            ops.gemm_group(specs)
        if need_x:
            dx = dxc if dxc.dtype == xdtype or xdtype not in (torch.float32, torch.bfloat16) else ops.cast(dxc, torch.empty(M, K, dtype=xdtype, device=dev))
            dx = dx.view(xshape)
        db = ops.colsum(dyc, torch.zeros(N, dtype=torch.float32, device=dev)) if (has_b or has_e) else None
        dres = None
        if has_res:
            rshape, rdtype = res_meta
            dres = dy.reshape(rshape)
            if dres.dtype != rdtype:
                dres = dres.to(rdtype)
        return (dx, None if dw is None else dw.to(wdtype), (db if has_b else None), (db if has_e else None), dres,
                None, None, None, None, None, None)


def linear(x, weight, bias=None, extra_bias=None, act=ops.ACT_NONE, dtype=torch.bfloat16, out_dtype=None, residual=None,
           drop_p=0.0, drop_seed=0, drop_site=0):
    if not x.is_cuda:
        raise RuntimeError("m3ae_b200.functional.linear needs CUDA tensors: there is no CPU fallback")
    return _LinearFn.apply(x, weight, bias, extra_bias, residual, act, dtype, out_dtype or dtype, float(drop_p), int(drop_seed),
                           int(drop_site))


class _LayerNormFn(torch.autograd.Function):
    """nn.LayerNorm (m3ae_decoder.py:46-49, clip_model.py:27-33, m3ae_module.py:122) on the C-ABI kernels; x [.., H] in fp32
    or bf16, affine fp32, statistics in fp32 (which is what CLIP's fp32-cast LayerNorm subclass enforces)."""

    @staticmethod
    def forward(ctx, x, weight, bias, eps):
        x2 = x.reshape(-1, x.shape[-1]).contiguous()
        y, mean, rstd = ops.layernorm_fwd(x2, weight.detach().float().contiguous(), bias.detach().float().contiguous(), eps)
        ctx.save_for_backward(x2, mean, rstd, weight)
        ctx.shape = x.shape
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, mean, rstd, weight = ctx.saved_tensors
        H = x2.shape[-1]
        d2 = dy.reshape(-1, H).contiguous()
        if d2.dtype != x2.dtype:
            d2 = ops.cast(d2, torch.empty_like(x2))
        dg = torch.zeros(H, dtype=torch.float32, device=x2.device)
        db = torch.zeros(H, dtype=torch.float32, device=x2.device)
        dx = ops.layernorm_bwd(d2, x2, mean, rstd, weight.detach().float().contiguous(), dg, db)
        return dx.view(ctx.shape), dg.to(weight.dtype), db.to(weight.dtype), None


def layernorm(x, weight, bias, eps):
    return _LayerNormFn.apply(x, weight, bias, eps)


class _AttentionFn(torch.autograd.Function):
    """Multi-head attention core on packed projections: q [B*Lq, H*dh], k/v [B*Lk, H*dh] (may be column slices of a fused
    projection buffer), additive key mask [B,Lk] or None, optional causal mask, optional dropout on the probabilities
    (nn.MultiheadAttention(dropout=p) at m3ae_decoder.py:42-43; HF RobertaSelfAttention).  Returns ctx [B*Lq, H*dh]."""

    @staticmethod
    def forward(ctx, q, k, v, mask, B, H, Lq, Lk, dh, causal, drop_p, drop_seed, drop_site):
        out = torch.empty(B * Lq, H * dh, dtype=q.dtype, device=q.device)
        lse = torch.empty(B, H, Lq, dtype=torch.float32, device=q.device)
        ops.attn_fwd(q, k, v, out, lse, B=B, H=H, Lq=Lq, Lk=Lk, dh=dh, mask=mask, causal=causal, drop_p=drop_p,
                     drop_seed=drop_seed, drop_stream=drop_site)
        ctx.save_for_backward(q, k, v, out, lse, mask)
        ctx.dims = (B, H, Lq, Lk, dh, causal, drop_p, drop_seed, drop_site)
        return out

    @staticmethod
    def backward(ctx, dout):
        q, k, v, out, lse, mask = ctx.saved_tensors
        B, H, Lq, Lk, dh, causal, drop_p, drop_seed, drop_site = ctx.dims
        dout = dout.contiguous()
        dq = torch.empty(B * Lq, H * dh, dtype=q.dtype, device=q.device)
        dk = torch.empty(B * Lk, H * dh, dtype=q.dtype, device=q.device)
        dv = torch.empty(B * Lk, H * dh, dtype=q.dtype, device=q.device)
        delta = torch.empty(B, H, Lq, dtype=torch.float32, device=q.device)
        ops.attn_bwd(q, k, v, out, lse, B=B, H=H, Lq=Lq, Lk=Lk, dh=dh, mask=mask, causal=causal, dout=dout, dq=dq, dk=dk,
                     dv=dv, delta=delta, drop_p=drop_p, drop_seed=drop_seed, drop_stream=drop_site)
        return (dq, dk, dv) + (None,) * 10


def attention(q, k, v, mask, B, H, Lq, Lk, dh, causal=False, drop_p=0.0, drop_seed=0, drop_site=0):
    return _AttentionFn.apply(q, k, v, mask, B, H, Lq, Lk, dh, causal, float(drop_p), int(drop_seed), int(drop_site))


class _FFNFn(torch.autograd.Function):
    """y = drop(act(x W0^T + b0) W2^T + b2) (+ residual): the decoder FFN (ReLU, m3ae_decoder.py:45,85-87), the CLIP MLP
    (QuickGELU, clip_model.py:46-50,62) and the RoBERTa FFN (erf GELU).  The activation's backward is fused into the dgrad
    epilogue of the second GEMM (needs the saved pre-activation)."""

    @staticmethod
    def forward(ctx, x, w0, b0, w2, b2, residual, act, dtype, drop_p, drop_seed, drop_site):
        M, K = x.shape
        F = w0.shape[0]
        dev = x.device
        xc = _as_compute(x, dtype)
        w0c = _as_compute(w0.detach(), dtype)
        w2c = _as_compute(w2.detach(), dtype)
        u = torch.empty(M, F, dtype=dtype, device=dev)
        h = torch.empty(M, F, dtype=dtype, device=dev)
        ops.gemm(xc, w0c, h, bias=b0.detach().float().contiguous(), act=act, aux_out=u)
        y = torch.empty(M, w2.shape[0], dtype=dtype, device=dev)
        rc = None
        if residual is not None:
            rc = _as_compute(residual, dtype)
        ops.gemm(h, w2c, y, bias=b2.detach().float().contiguous(), residual=rc, drop_p=drop_p, drop_seed=drop_seed, drop_stream=drop_site)
        ctx.save_for_backward(xc, w0c, w2c, u, h)
        ctx.meta = (act, dtype, residual is not None, x.dtype, w0.dtype, None if residual is None else residual.dtype,
                    drop_p, drop_seed, drop_site)
        return y

    @staticmethod
    def backward(ctx, dy):
        xc, w0c, w2c, u, h = ctx.saved_tensors
        act, dtype, has_res, xdtype, wdtype, rdtype, drop_p, drop_seed, drop_site = ctx.meta
        M, K = xc.shape
        F = w0c.shape[0]
        N = w2c.shape[0]
        dev = dy.device
        dyc = _as_compute(dy, dtype)
        dres = None
        if has_res:
            dres = dy if dy.dtype == rdtype else dy.to(rdtype)
        if drop_p > 0.0:
            dyc = ops.dropout_apply(dyc, torch.empty_like(dyc), drop_p, drop_seed, drop_site)
        du = torch.empty(M, F, dtype=dtype, device=dev)
        dw2 = torch.zeros(N, F, dtype=torch.float32, device=dev)
        MN = ops.MAJOR_MN
        ops.gemm_group([ops.GemmSpec(dyc, w2c, du, b_major=MN, aux_in=u, dact=act),
                        ops.GemmSpec(dyc, h, dw2, a_major=MN, b_major=MN, accumulate=True, split_k=_wgrad_split(M, N, F))])
        dx = torch.empty(M, K, dtype=dtype, device=dev)
        dw0 = torch.zeros(F, K, dtype=torch.float32, device=dev)
        ops.gemm_group([ops.GemmSpec(du, w0c, dx, b_major=MN),
                        ops.GemmSpec(du, xc, dw0, a_major=MN, b_major=MN, accumulate=True, split_k=_wgrad_split(M, F, K))])
        db2 = ops.colsum(dyc, torch.zeros(N, dtype=torch.float32, device=dev))
        db0 = ops.colsum(du, torch.zeros(F, dtype=torch.float32, device=dev))
        if dx.dtype != xdtype:
            dx = ops.cast(dx, torch.empty(M, K, dtype=xdtype, device=dev))
        return dx, dw0.to(wdtype), db0, dw2.to(wdtype), db2, dres, None, None, None, None, None


def ffn(x, w0, b0, w2, b2, residual=None, act=ops.ACT_RELU, dtype=torch.bfloat16, drop_p=0.0, drop_seed=0, drop_site=0):
    return _FFNFn.apply(x, w0, b0, w2, b2, residual, act, dtype, float(drop_p), int(drop_seed), int(drop_site))


class _DropoutFn(torch.autograd.Function):
    """nn.Dropout on a 2-D activation (decoder embedding dropout, m3ae_decoder.py:128; RoBERTa embedding dropout):
    counter-based Philox mask keyed by (seed, site, element), regenerated in backward."""

    @staticmethod
    def forward(ctx, x, p, seed, site):
        x2 = x.reshape(-1, x.shape[-1]).contiguous()
        ctx.meta = (x.shape, p, seed, site)
        return ops.dropout_apply(x2, torch.empty_like(x2), p, seed, site).view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        shape, p, seed, site = ctx.meta
        d2 = dy.reshape(-1, shape[-1]).contiguous()
        return ops.dropout_apply(d2, torch.empty_like(d2), p, seed, site).view(shape), None, None, None


def dropout(x, p, seed, site=0):
    if p <= 0.0:
        return x
    return _DropoutFn.apply(x, float(p), int(seed), int(site))


class _ActivationFn(torch.autograd.Function):
    """y = act(x) element-wise: Pooler tanh (prediction_heads.py:18), vqa_head GELU (m3ae_module.py:123)."""

    @staticmethod
    def forward(ctx, x, act):
        xc = x.contiguous()
        ctx.save_for_backward(xc)
        ctx.act = act
        return ops.activation_fwd(xc, act)

    @staticmethod
    def backward(ctx, dy):
        (xc,) = ctx.saved_tensors
        d = dy.contiguous()
        if d.dtype != xc.dtype:
            d = d.to(xc.dtype)
        return ops.activation_bwd(d, xc, ctx.act), None


def activation(x, act):
    return _ActivationFn.apply(x, int(act))


class _BCEWithLogitsFn(torch.autograd.Function):
    """objectives.py:201: F.binary_cross_entropy_with_logits(logits, targets) * targets.shape[1] — loss and dlogits in
    one pass over the [B, labels] logits."""

    @staticmethod
    def forward(ctx, logits, targets):
        lg = logits.float().contiguous()
        tg = targets.float().contiguous()
        B = lg.shape[0]
        loss, dx = ops.bce_with_logits(lg, tg, 1.0 / B, 1.0 / B, want_grad=ctx.needs_input_grad[0])
        ctx.save_for_backward(dx)
        ctx.in_dtype = logits.dtype
        return loss

    @staticmethod
    def backward(ctx, dloss):
        (dx,) = ctx.saved_tensors
        return (dx * dloss).to(ctx.in_dtype), None


def vqa_bce_loss(logits, targets):
    return _BCEWithLogitsFn.apply(logits, targets)


class _CrossEntropyFn(torch.autograd.Function):
    """m3ae_decoder.py:365-368: CrossEntropyLoss(ignore_index=pad)(logits^T, golden) — the mean over the non-pad golden
    tokens (the `* padding_mask ... / padding_mask.sum()` that follows multiplies and divides a scalar by the same sum)."""

    @staticmethod
    def forward(ctx, logits, target, ignore_index):
        V = logits.shape[-1]
        lg = logits.reshape(-1, V).float()
        if lg.stride(1) != 1:
            lg = lg.contiguous()
        tg = target.reshape(-1).contiguous()
        n = (tg != ignore_index).sum().clamp_min(1).float()
        w = (1.0 / n).expand(tg.shape[0]).contiguous()
        loss, dx = ops.cross_entropy(lg, tg, w, ignore_index, want_grad=ctx.needs_input_grad[0])
        ctx.save_for_backward(dx)
        ctx.meta = (logits.shape, logits.dtype)
        return loss

    @staticmethod
    def backward(ctx, dloss):
        (dx,) = ctx.saved_tensors
        shape, dtype = ctx.meta
        return (dx * dloss).view(shape).to(dtype), None, None


def cross_entropy(logits, target, ignore_index=-100):
    return _CrossEntropyFn.apply(logits, target, int(ignore_index))
