"""Tensor-level wrappers over the C-ABI (libm3ae_b200.so).  torch is used for device memory and streams only.

Every function here launches hand-written sm_100a kernels; none has a PyTorch fallback.
"""
import ctypes as C

import torch

from . import _lib as L
from ._lib import ACT_GELU, ACT_NONE, ACT_QUICKGELU, ACT_RELU, ACT_TANH, ATTN_MMA_SYNC, ATTN_NARROW, BF16, F32, GEMM_AUTO, GEMM_SIMT, GEMM_TCGEN05, MAJOR_K, MAJOR_MN  # noqa: F401


def _code(dtype):
    if dtype == torch.float32:
        return F32
    if dtype == torch.bfloat16:
        return BF16
    raise TypeError(f"unsupported dtype {dtype}: the M3AE B200 path takes float32 or bfloat16")


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)
_raw_device = getattr(torch._C, "_cuda_getDevice", None)


def _stream():
    """cudaStream_t of torch's current stream on the current device.  torch.cuda.current_stream().cuda_stream builds a Stream
    object through several Python layers (~17 us per call: with ~860 library launches per whole-model step that was 40 % of the
    host time, tools/whole_step_cprofile.py); the raw-handle call the compiler back ends use costs well under a microsecond."""
    if _raw_stream is not None and _raw_device is not None:
        return C.c_void_p(_raw_stream(_raw_device()))
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("m3ae_b200 ops need CUDA tensors (sm_100a); there is no CPU path")


def launch_count():
    return int(L.load().m3ae_launch_count())


def set_attention_backend(backend):
    L.check(L.load().m3ae_set_attention_backend(int(backend)), "m3ae_set_attention_backend")


def set_gemm_epilogue(generic):
    """Test hook: True forces the generic per-element epilogue of the tcgen05 GEMM."""
    L.check(L.load().m3ae_set_gemm_epilogue(1 if generic else 0), "m3ae_set_gemm_epilogue")


def set_layernorm_bwd_staging(registers_only):
    """Test hook: True = LayerNorm backward without the bulk-copy staging ring."""
    L.check(L.load().m3ae_set_layernorm_bwd_staging(1 if registers_only else 0), "m3ae_set_layernorm_bwd_staging")


def set_gemm_cta_pair(mode):
    """0 auto, 1 one CTA per 128x256 tile, 2 CTA pairs (tcgen05.mma.cta_group::2, 256x256 tiles), 3 clusters of two pairs
    that share one operand by TMA multicast (what auto picks for M > 128)."""
    L.check(L.load().m3ae_set_gemm_cta_pair(int(mode)), "m3ae_set_gemm_cta_pair")


class GemmSpec:
    """One problem of a grouped launch: out (+)= epi(alpha * A[M,K] · B[N,K]^T).

    a: [M,K] (a_major=K) or [K,M] (a_major=MN) 2-D tensor with unit inner stride; likewise b with N.
    """

    def __init__(self, a, b, out, *, a_major=MAJOR_K, b_major=MAJOR_K, bias=None, act=ACT_NONE, aux_out=None,
                 aux_in=None, dact=ACT_NONE, residual=None, drop_p=0.0, drop_seed=0, drop_stream=0, accumulate=False,
                 split_k=1, alpha=1.0, aux_is_grad=False, drop_bits=None, drop_bits_ready=False, amax=None):
        """amax = (vals fp32 [ceil(N/64), M], idx int32 [ceil(N/64), M]): the epilogue also records, per row and 64-column
        group, the maximum of the result and its column (reduced by argmax_partials); `out` may then be None."""
        _need_cuda(a, b, out, bias, aux_out, aux_in, residual, drop_bits)
        assert a.dim() == 2 and b.dim() == 2 and a.stride(1) == 1 and b.stride(1) == 1
        assert a.dtype == b.dtype
        if out is None:
            if amax is None:
                raise ValueError("GEMM needs an output tensor (or amax partials)")
        else:
            assert out.dim() == 2 and out.stride(1) == 1
        M, K = (a.shape if a_major == MAJOR_K else (a.shape[1], a.shape[0]))
        N, Kb = (b.shape if b_major == MAJOR_K else (b.shape[1], b.shape[0]))
        if K != Kb:
            raise ValueError(f"GEMM contraction mismatch: A gives K={K}, B gives K={Kb}")
        if out is not None and tuple(out.shape) != (M, N):
            raise ValueError(f"GEMM output shape {tuple(out.shape)} != ({M},{N})")
        if bias is not None:
            assert bias.dtype == torch.float32 and bias.numel() == N and bias.is_contiguous()
        for t in (aux_out, aux_in, residual):
            if t is not None:
                assert tuple(t.shape) == (M, N) and t.stride(1) == 1
        if aux_out is not None:
            assert aux_out.dtype == out.dtype and aux_out.stride(0) == out.stride(0)
        self.in_dtype = a.dtype
        if drop_bits is not None:  # keep bits of the epilogue dropout, [M, N/32] uint32 words (read by layernorm_bwd)
            assert N % 32 == 0 and drop_bits.is_contiguous() and drop_bits.numel() * drop_bits.element_size() == M * (N // 32) * 4
        self.keep = (a, b, out, bias, aux_out, aux_in, residual, drop_bits, amax)  # keep tensors alive until launch returns
        p = L.GemmProblem()
        p.a, p.b = a.data_ptr(), b.data_ptr()
        p.lda, p.ldb = a.stride(0), b.stride(0)
        p.a_major, p.b_major = a_major, b_major
        p.M, p.N, p.K = M, N, K
        p.split_k = split_k
        e = p.epi
        if out is not None:
            e.out, e.out_dtype, e.ldo = out.data_ptr(), _code(out.dtype), out.stride(0)
        else:
            e.out, e.out_dtype, e.ldo = None, F32, N
        if amax is not None:
            av, ai = amax
            _need_cuda(av, ai)
            G = (N + 63) // 64
            assert av.dtype == torch.float32 and ai.dtype == torch.int32 and av.is_contiguous() and ai.is_contiguous()
            assert tuple(av.shape) == (G, M) and tuple(ai.shape) == (G, M)
            e.amax_val, e.amax_idx = av.data_ptr(), ai.data_ptr()
        e.bias = bias.data_ptr() if bias is not None else None
        e.act = act
        e.aux_out = aux_out.data_ptr() if aux_out is not None else None
        if aux_in is not None:
            e.aux_in, e.aux_in_dtype, e.ld_aux_in = aux_in.data_ptr(), _code(aux_in.dtype), aux_in.stride(0)
        e.dact = dact
        if residual is not None:
            e.residual, e.res_dtype, e.ldr = residual.data_ptr(), _code(residual.dtype), residual.stride(0)
        e.drop_p, e.drop_seed, e.drop_stream = float(drop_p), int(drop_seed), int(drop_stream)
        e.accumulate = 1 if accumulate else 0
        e.alpha = float(alpha)
        e.aux_is_grad = 1 if aux_is_grad else 0
        e.drop_bits = drop_bits.data_ptr() if drop_bits is not None else None
        e.drop_bits_ready = 1 if (drop_bits is not None and drop_bits_ready) else 0
        self.problem = p


def gemm_group(specs, backend=GEMM_AUTO):
    """Run 1..4 independent GEMMs in one persistent launch."""
    n = len(specs)
    arr = (L.GemmProblem * n)(*[s.problem for s in specs])
    in_dtype = specs[0].in_dtype
    assert all(s.in_dtype == in_dtype for s in specs)
    L.check(L.load().m3ae_gemm(arr, n, _code(in_dtype), backend, _stream()), "m3ae_gemm")


def gemm(a, b, out, backend=GEMM_AUTO, **kw):
    gemm_group([GemmSpec(a, b, out, **kw)], backend)
    return out


def attn_params(q, k, v, out, lse, *, B, H, Lq, Lk, dh, mask=None, probs=None, scale=None, causal=False, drop_p=0.0,
                drop_seed=0, drop_stream=0, dout=None, dq=None, dk=None, dv=None, delta=None, drop_mask=None,
                drop_mask_ready=False):
    _need_cuda(q, k, v, out, lse, mask, probs, dout, dq, dk, dv, delta, drop_mask)
    for t in (q, k, v, out):
        assert t.dim() == 2 and t.stride(1) == 1 and t.dtype == q.dtype
    assert q.shape[0] == B * Lq and k.shape[0] == B * Lk and v.shape[0] == B * Lk and out.shape[0] == B * Lq
    assert lse.dtype == torch.float32 and lse.numel() == B * H * Lq and lse.is_contiguous()
    if mask is not None:
        assert mask.dtype == torch.float32 and mask.numel() == B * Lk and mask.is_contiguous()
    if probs is not None:
        assert probs.dtype == torch.float32 and probs.numel() == B * H * Lq * Lk and probs.is_contiguous()
    p = L.AttnParams()
    p.q, p.k, p.v = q.data_ptr(), k.data_ptr(), v.data_ptr()
    p.ldq, p.ldk, p.ldv = q.stride(0), k.stride(0), v.stride(0)
    p.mask = mask.data_ptr() if mask is not None else None
    p.out, p.ldo = out.data_ptr(), out.stride(0)
    p.lse = lse.data_ptr()
    p.probs = probs.data_ptr() if probs is not None else None
    p.B, p.H, p.Lq, p.Lk, p.dh = B, H, Lq, Lk, dh
    p.scale = float(scale if scale is not None else dh ** -0.5)
    p.causal = 1 if causal else 0
    p.drop_p, p.drop_seed, p.drop_stream = float(drop_p), int(drop_seed), int(drop_stream)
    p.dtype = _code(q.dtype)
    if drop_mask is not None:
        assert drop_mask.dtype == torch.uint8 and drop_mask.is_contiguous() and drop_mask.numel() == B * H * Lq * drop_mask_ld(Lk)
        p.drop_mask = drop_mask.data_ptr()
        p.drop_mask_ready = 1 if drop_mask_ready else 0
    if dout is not None:
        for t in (dout, dq, dk, dv):
            assert t.dim() == 2 and t.stride(1) == 1 and t.dtype == q.dtype
        assert dout.stride(0) == out.stride(0)
        assert delta.dtype == torch.float32 and delta.numel() == B * H * Lq
        p.dout, p.dq, p.dk, p.dv = dout.data_ptr(), dq.data_ptr(), dk.data_ptr(), dv.data_ptr()
        p.lddq, p.lddk, p.lddv = dq.stride(0), dk.stride(0), dv.stride(0)
        p.delta = delta.data_ptr()
    return p


def drop_mask_ld(Lk):
    """bytes per attention row of the dropout keep-bit cache (64-key tiles, 8 bytes each)."""
    return 8 * ((Lk + 63) // 64)


def attn_fwd(q, k, v, out, lse, **kw):
    p = attn_params(q, k, v, out, lse, **kw)
    L.check(L.load().m3ae_attn_fwd(C.byref(p), _stream()), "m3ae_attn_fwd")
    return out


def attn_bwd(q, k, v, out, lse, **kw):
    p = attn_params(q, k, v, out, lse, **kw)
    L.check(L.load().m3ae_attn_bwd(C.byref(p), _stream()), "m3ae_attn_bwd")


def attn_probs(q, k, lse, probs, *, B, H, Lq, Lk, dh, mask=None, scale=None, causal=False, drop_p=0.0, drop_seed=0, drop_stream=0):
    """probs [B,H,Lq,Lk] fp32 = dropout(exp(scale*q.k^T + mask - lse)) from a forward call's q, k and saved lse."""
    out = q  # attn_params wants an `out` of q's geometry; the probs kernel never touches it
    p = attn_params(q, k, k, out, lse, B=B, H=H, Lq=Lq, Lk=Lk, dh=dh, mask=mask, probs=probs, scale=scale, causal=causal,
                    drop_p=drop_p, drop_seed=drop_seed, drop_stream=drop_stream)
    L.check(L.load().m3ae_attn_probs(C.byref(p), _stream()), "m3ae_attn_probs")
    return probs


def activation_fwd(x, act, y=None):
    _need_cuda(x)
    assert x.is_contiguous()
    y = torch.empty_like(x) if y is None else y
    L.check(L.load().m3ae_activation_fwd(_ptr(x), _ptr(y), x.numel(), int(act), _code(x.dtype), _stream()), "m3ae_activation_fwd")
    return y


def activation_bwd(dy, x, act, dx=None):
    _need_cuda(dy, x)
    assert x.is_contiguous() and dy.is_contiguous() and dy.dtype == x.dtype and dy.numel() == x.numel()
    dx = torch.empty_like(x) if dx is None else dx
    L.check(L.load().m3ae_activation_bwd(_ptr(dy), _ptr(x), _ptr(dx), x.numel(), int(act), _code(x.dtype), _stream()),
            "m3ae_activation_bwd")
    return dx


def bce_with_logits(logits, targets, scale_loss, scale_grad, want_grad=True):
    """-> (loss scalar tensor, dlogits or None); loss = scale_loss * sum bce(logits, targets)."""
    _need_cuda(logits, targets)
    assert logits.dtype == torch.float32 and targets.dtype == torch.float32 and logits.is_contiguous() and targets.is_contiguous()
    assert logits.shape == targets.shape
    loss = torch.empty((), dtype=torch.float32, device=logits.device)
    dx = torch.empty_like(logits) if want_grad else None
    L.check(L.load().m3ae_bce_with_logits(_ptr(logits), _ptr(targets), _ptr(loss), _ptr(dx), logits.numel(), float(scale_loss),
                                          float(scale_grad), _stream()), "m3ae_bce_with_logits")
    return loss, dx


def cross_entropy(logits, target, row_weight, ignore_index, want_grad=True):
    """logits [M,V] fp32, target [M] int64, row_weight [M] fp32 -> (loss scalar, dlogits [M,V] or None)."""
    _need_cuda(logits, target, row_weight)
    assert logits.dim() == 2 and logits.dtype == torch.float32 and logits.stride(1) == 1
    assert target.dtype == torch.int64 and target.is_contiguous() and row_weight.dtype == torch.float32 and row_weight.is_contiguous()
    M, V = logits.shape
    loss = torch.empty((), dtype=torch.float32, device=logits.device)
    dx = torch.empty(M, V, dtype=torch.float32, device=logits.device) if want_grad else None
    L.check(L.load().m3ae_cross_entropy(_ptr(logits), logits.stride(0), _ptr(target), _ptr(row_weight), _ptr(loss), _ptr(dx), V, M, V,
                                        int(ignore_index), _stream()), "m3ae_cross_entropy")
    return loss, dx


def layernorm_fwd(x, gamma, beta, eps, y=None, mean=None, rstd=None):
    _need_cuda(x, gamma, beta)
    assert x.is_contiguous() and gamma.dtype == torch.float32 and beta.dtype == torch.float32
    H = x.shape[-1]
    M = x.numel() // H
    y = torch.empty_like(x) if y is None else y
    mean = torch.empty(M, dtype=torch.float32, device=x.device) if mean is None else mean
    rstd = torch.empty(M, dtype=torch.float32, device=x.device) if rstd is None else rstd
    L.check(L.load().m3ae_layernorm_fwd(_ptr(x), _ptr(gamma), _ptr(beta), _ptr(y), _ptr(mean), _ptr(rstd), M, H,
                                        float(eps), _code(x.dtype), _stream()), "m3ae_layernorm_fwd")
    return y, mean, rstd


def layernorm_bwd(dy, x, mean, rstd, gamma, dgamma=None, dbeta=None, dcolsum=None, dx=None, drop_p=0.0, drop_seed=0,
                  drop_stream=0, drop_bits=None):
    """dgamma / dbeta / dcolsum are accumulated into (+=); pass zeroed fp32 buffers for plain gradients.
    Returns dx, or (dx, dx_drop) when drop_p > 0 (dx_drop = dx pushed back through the producing GEMM's dropout)."""
    _need_cuda(dy, x, mean, rstd, gamma, dgamma, dbeta, dcolsum)
    assert dy.is_contiguous() and x.is_contiguous() and dy.dtype == x.dtype
    H = x.shape[-1]
    M = x.numel() // H
    dx = torch.empty_like(x) if dx is None else dx
    dx_drop = torch.empty_like(x) if drop_p > 0 else None
    L.check(L.load().m3ae_layernorm_bwd(_ptr(dy), _ptr(x), _ptr(mean), _ptr(rstd), _ptr(gamma), _ptr(dx), _ptr(dgamma),
                                        _ptr(dbeta), _ptr(dcolsum), M, H, _code(x.dtype), _ptr(dx_drop), float(drop_p),
                                        int(drop_seed), int(drop_stream), _ptr(drop_bits), _stream()),
            "m3ae_layernorm_bwd")
    return (dx, dx_drop) if drop_p > 0 else dx


def colsum(x, out):
    """out[N] (fp32) += column sums of the 2-D tensor x[M,N]."""
    _need_cuda(x, out)
    assert x.dim() == 2 and x.stride(1) == 1 and out.dtype == torch.float32 and out.numel() == x.shape[1]
    L.check(L.load().m3ae_colsum(_ptr(x), _ptr(out), x.shape[0], x.shape[1], x.stride(0), _code(x.dtype), _stream()),
            "m3ae_colsum")
    return out


def cast(src, dst):
    _need_cuda(src, dst)
    assert src.is_contiguous() and dst.is_contiguous() and src.numel() == dst.numel()
    L.check(L.load().m3ae_cast(_ptr(src), _code(src.dtype), _ptr(dst), _code(dst.dtype), src.numel(), _stream()),
            "m3ae_cast")
    return dst


class CastGroup:
    """A cached list of (src, dst) casts issued as ONE launch (m3ae_cast_group)."""

    def __init__(self, pairs):
        assert 1 <= len(pairs) <= L.MAX_CAST_GROUP
        for src, dst in pairs:
            _need_cuda(src, dst)
            assert src.is_contiguous() and dst.is_contiguous() and src.numel() == dst.numel()
        self.keep = pairs
        self.n = len(pairs)
        self.arr = (L.CastItem * self.n)(*[L.CastItem(s.data_ptr(), d.data_ptr(), s.numel(), _code(s.dtype), _code(d.dtype)) for s, d in pairs])
        self.fn = L.load().m3ae_cast_group

    def run(self):
        L.check(self.fn(self.arr, self.n, _stream()), "m3ae_cast_group")


class DropoutBits:
    """A cached grouped launch of m3ae_dropout_bits: keep bits of up to 12 dropout sites (attention caches [B*H*Lq, 8*ceil(Lk/64)]
    bytes, hidden caches [M, N/32] uint32) generated in one streaming kernel ahead of the kernels that apply them.
    items: (kind 0 attention | 1 hidden, out tensor, rows, n, ld, p, site); seeds are set per call."""

    def __init__(self, items):
        assert 1 <= len(items) <= L.MAX_DROPBITS
        self.keep = [it[1] for it in items]
        self.n = len(items)
        self.arr = (L.DropBitsItem * self.n)()
        for a, (kind, out, rows, n, ld, p, site) in zip(self.arr, items):
            _need_cuda(out)
            assert out.is_contiguous()
            need = rows * (8 * ((n + 63) // 64) if kind == 0 else (n // 32) * 4)
            assert out.numel() * out.element_size() == need, "keep-bit buffer has the wrong size"
            a.out, a.rows, a.ld, a.n, a.kind, a.p, a.site = out.data_ptr(), rows, ld, n, kind, float(p), int(site)
        self.fn = L.load().m3ae_dropout_bits

    def run(self, seeds):
        for a, s in zip(self.arr, seeds):
            a.seed = int(s)
        L.check(self.fn(self.arr, self.n, _stream()), "m3ae_dropout_bits")


def add_rowvec(a, b, out):
    _need_cuda(a, b, out)
    assert a.is_contiguous() and out.is_contiguous() and b.dtype == torch.float32 and b.is_contiguous()
    N = a.shape[-1]
    L.check(L.load().m3ae_add_rowvec(_ptr(a), _ptr(b), _ptr(out), a.numel() // N, N, _code(a.dtype), _code(out.dtype),
                                     _stream()), "m3ae_add_rowvec")
    return out


def dropout_apply(x, y, p, seed, stream_id):
    _need_cuda(x, y)
    assert x.dim() == 2 and x.stride(1) == 1 and y.stride() == x.stride() and x.dtype == y.dtype
    L.check(L.load().m3ae_dropout_apply(_ptr(x), _ptr(y), x.shape[0], x.shape[1], x.stride(0), float(p), int(seed),
                                        int(stream_id), _code(x.dtype), _stream()), "m3ae_dropout_apply")
    return y


def decode_attn(q, kcache, vcache, out, *, H, length, dh, scale=None):
    """q/out: [B, H*dh]; caches: [B, Lmax, H*dh]."""
    _need_cuda(q, kcache, vcache, out)
    B = q.shape[0]
    assert kcache.dim() == 3 and kcache.stride(2) == 1 and vcache.stride(2) == 1
    L.check(L.load().m3ae_decode_attn(_ptr(q), q.stride(0), _ptr(kcache), _ptr(vcache), kcache.stride(1), vcache.stride(1),
                                      kcache.stride(0), vcache.stride(0), _ptr(out), out.stride(0), B, H, int(length), dh,
                                      float(scale if scale is not None else dh ** -0.5), _code(q.dtype), _stream()),
            "m3ae_decode_attn")
    return out


def argmax_partials(vals, idx, out=None):
    """Row argmax (int64 [M]) from the per-64-column partials a GEMM epilogue wrote (GemmSpec(amax=...)); first maximum wins."""
    _need_cuda(vals, idx)
    G, M = vals.shape
    out = torch.empty(M, dtype=torch.int64, device=vals.device) if out is None else out
    assert out.dtype == torch.int64 and out.is_contiguous() and out.numel() == M
    L.check(L.load().m3ae_argmax_partials(_ptr(vals), _ptr(idx), G, M, _ptr(out), _stream()), "m3ae_argmax_partials")
    return out


def set_pdl(on):
    """Programmatic dependent launch of the library's kernels on / off (default on)."""
    L.check(L.load().m3ae_set_pdl(1 if on else 0), "m3ae_set_pdl")


def set_gemm_reserved_sms(n):
    """SMs the persistent tcgen05 GEMM leaves free for kernels of other streams (the NCCL all-reduce under data parallelism)."""
    L.check(L.load().m3ae_set_gemm_reserved_sms(int(n)), "m3ae_set_gemm_reserved_sms")


def set_gemm_skinny(mode):
    """Test hook for single-CTA GEMM launches (every problem M <= 128): 0 auto, 1 always 128x256 tiles, 2 always 128x64 tiles."""
    L.check(L.load().m3ae_set_gemm_skinny(int(mode)), "m3ae_set_gemm_skinny")


def argmax_rows(logits, out=None):
    _need_cuda(logits)
    assert logits.dim() == 2 and logits.stride(1) == 1
    out = torch.empty(logits.shape[0], dtype=torch.int64, device=logits.device) if out is None else out
    L.check(L.load().m3ae_argmax_rows(_ptr(logits), logits.stride(0), _ptr(out), logits.shape[0], logits.shape[1],
                                      _code(logits.dtype), _stream()), "m3ae_argmax_rows")
    return out


def decoder_embed(ids, pos, table, pe, out):
    _need_cuda(ids, pos, table, pe, out)
    assert ids.dtype == torch.int64 and pos.dtype == torch.int32 and table.dtype == torch.float32 and pe.dtype == torch.float32
    assert table.is_contiguous() and pe.is_contiguous() and out.is_contiguous()
    D = table.shape[1]
    L.check(L.load().m3ae_decoder_embed(_ptr(ids), _ptr(pos), _ptr(table), _ptr(pe), _ptr(out), ids.numel(), D,
                                        _code(out.dtype), _stream()), "m3ae_decoder_embed")
    return out


class SkinnyLinear:
    """A cached m3ae_skinny_linear launch (decoder token step): out = act(LN?(x) w^T + bias) (+ residual), M <= 128 rows.
    `outs`: one output [M, N] or up to three [M, N/3] segments (column blocks of the fused projection, each with its own
    row pitch — e.g. q and the K / V cache rows of the current position).  `ln` = (gamma, beta, eps) folds the LayerNorm
    in front of the projection into the activation load (K = 768).  K > 768 is split across CTAs through `workspace`
    (zeroed fp32 [128, N]) and `counters` (zeroed int32), which the kernel leaves zeroed."""

    def __init__(self, x, w, outs, bias=None, residual=None, act=ACT_NONE, ln=None, workspace=None, counters=None):
        _need_cuda(x, w)
        outs = list(outs) if isinstance(outs, (list, tuple)) else [outs]
        M, K = x.shape
        N = w.shape[0]
        assert x.dtype == w.dtype == torch.bfloat16 and w.shape[1] == K and x.stride(1) == 1 and w.stride(1) == 1
        assert N % len(outs) == 0 and all(o.dtype == torch.bfloat16 and o.shape == (M, N // len(outs)) and o.stride(1) == 1 for o in outs)
        p = L.SkinnyParams()
        p.x, p.ldx, p.w, p.ldw = _ptr(x), x.stride(0), _ptr(w), w.stride(0)
        p.bias = _ptr(bias) if bias is not None else None
        if bias is not None:
            assert bias.dtype == torch.float32 and bias.is_contiguous() and bias.numel() == N
        if residual is not None:
            assert residual.dtype == torch.bfloat16 and residual.shape == (M, N) and residual.stride(1) == 1
            p.residual, p.ldr = _ptr(residual), residual.stride(0)
        for i, o in enumerate(outs):
            p.out[i], p.ldo[i] = _ptr(o), o.stride(0)
        p.seg_cols, p.M, p.N, p.K, p.act = N // len(outs), M, N, K, int(act)
        p.ksplit = 1
        if K > 768:
            p.ksplit = next(s for s in range(2, 65) if K % s == 0 and K // s <= 768 and (K // s) % 16 == 0 and (K // s) <= 512)
            assert workspace is not None and counters is not None and workspace.dtype == torch.float32 and workspace.numel() >= 128 * N
            assert counters.dtype == torch.int32 and counters.numel() >= 256
            p.ws, p.counters = _ptr(workspace), _ptr(counters)
        if ln is not None:
            gamma, beta, eps = ln
            assert gamma.dtype == beta.dtype == torch.float32 and gamma.is_contiguous() and beta.is_contiguous() and K == 768
            p.ln_gamma, p.ln_beta, p.ln_eps = _ptr(gamma), _ptr(beta), float(eps)
        self.p, self.keep = p, (x, w, outs, bias, residual, ln, workspace, counters)
        self.fn = L.load().m3ae_skinny_linear

    def run(self):
        L.check(self.fn(C.byref(self.p), _stream()), "m3ae_skinny_linear")


def decoder_next_token(x, xn, table, pe, pos, ids_in=None, ids_out=None, amax=None, ln=None):
    """One launch for the greedy decoder's token hand-over: (optional) row argmax from the vocabulary projection's partials
    `amax` = (values [G, M] fp32, columns [G, M] int32) -> ids_out, x = 2*E[id] + pe[pos], xn = LayerNorm(x) (ln = (gamma, beta,
    eps); xn None skips it)."""
    M, D = x.shape
    assert x.dtype == torch.bfloat16 and x.is_contiguous() and table.dtype == torch.float32 and pe.dtype == torch.float32
    gamma, beta, eps = ln if ln is not None else (None, None, 0.0)
    val, idx, G = (amax[0], amax[1], amax[0].shape[0]) if amax is not None else (None, None, 0)
    L.check(L.load().m3ae_decoder_next_token(_ptr(val) if val is not None else None, _ptr(idx) if idx is not None else None, G, M,
                                             _ptr(ids_in) if ids_in is not None else None, _ptr(ids_out) if ids_out is not None else None,
                                             int(pos), _ptr(table), _ptr(pe), _ptr(gamma) if gamma is not None else None,
                                             _ptr(beta) if beta is not None else None, float(eps), _ptr(x),
                                             _ptr(xn) if xn is not None else None, D, _stream()), "m3ae_decoder_next_token")
