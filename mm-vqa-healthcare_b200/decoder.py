"""Drop-in for the reference's generative answer head `Decoder` (m3ae/modules/m3ae_decoder.py:93-182) on the sm_100a
kernels: same constructor, same state_dict keys (`dec_layers.{i}.{mha1,mha2}.{in_proj_weight,in_proj_bias,out_proj.*}`,
`dec_layers.{i}.ffn.{0,2}.*`, `dec_layers.{i}.{pre_norm,layernorm1,layernorm2,layernorm3}.*`,
`target_embedding.weight`, `positional_encoding.pe`, `final_linear.*`), same `forward` / `search_path` signatures.

Reference quirks that are reproduced because token IDs must match:
  * embedding is 2*E[ids] + pe (`target_embed += positional_encoding(target_embed)`, :125-127 with :36);
  * every layer is fed the embedding and only the LAST layer's output survives (:131-134), so
    logits == final_linear(dec_layers[-1](embed, memory)); layers 0..n-2 are dead compute and receive no gradient.
    They are skipped unless `return_all_attentions=True` asks for their attention maps;
  * DecoderLayer mixes pre- and post-norm: residuals are taken BEFORE layernorm1 / layernorm2 (:54-90);
  * search_path re-runs the whole prefix each step without a cache and stops only when all rows hit [SEP]; here the
    same tokens come from a KV-cached single-token step: cross-attention K/V of the memory are projected once, the
    self-attention K/V of each new token are appended to a cache by the projection GEMM itself.

Per greedy step at B=128 (bf16): live-layer weights 18.9 MB + vocab projection 46.9 MB + cross K/V cache 141 MB are
streamed once -> HBM-bound.
"""
import math

import torch
import torch.nn as nn

from . import functional as Fn
from . import ops


class PositionalEncoding(nn.Module):
    """m3ae_decoder.py:22-36 (buffer `pe` [1, max_len, d_model])."""

    def __init__(self, d_model, max_len=1024):
        super().__init__()
        pe = torch.zeros(max_len, d_model)
        position = torch.arange(0, max_len, dtype=torch.float).unsqueeze(1)
        div_term = torch.exp(torch.arange(0, d_model, 2).float() * (-math.log(10000.0) / d_model))
        pe[:, 0::2] = torch.sin(position * div_term)
        pe[:, 1::2] = torch.cos(position * div_term)
        self.register_buffer("pe", pe.unsqueeze(0))


class DecoderLayer(nn.Module):
    """Parameter container with the reference's names (m3ae_decoder.py:38-52)."""

    def __init__(self, d_model, num_heads, d_ff, dropout=0.1):
        super().__init__()
        self.mha1 = nn.MultiheadAttention(embed_dim=d_model, num_heads=num_heads, dropout=dropout, batch_first=True)
        self.mha2 = nn.MultiheadAttention(embed_dim=d_model, num_heads=num_heads, dropout=dropout, batch_first=True)
        self.ffn = nn.Sequential(nn.Linear(d_model, d_ff), nn.ReLU(), nn.Linear(d_ff, d_model))
        self.identity = nn.Identity()
        self.pre_norm = nn.LayerNorm(d_model)
        self.layernorm1 = nn.LayerNorm(d_model)
        self.layernorm2 = nn.LayerNorm(d_model)
        self.layernorm3 = nn.LayerNorm(d_model)
        self.dropout1 = nn.Dropout(dropout)
        self.dropout2 = nn.Dropout(dropout)
        self.dropout3 = nn.Dropout(dropout)


class _LazyMaps(dict):
    """dict of attention maps whose values are computed on first access (keys are all present from the start)."""

    def defer(self, key, fn):
        dict.__setitem__(self, key, None)
        self.__dict__.setdefault("_fns", {})[key] = fn

    def __getitem__(self, key):
        v = dict.__getitem__(self, key)
        if v is None and key in self.__dict__.get("_fns", {}):
            v = self._fns.pop(key)()
            dict.__setitem__(self, key, v)
        return v

    def get(self, key, default=None):
        return self[key] if key in self else default

    def materialize(self):
        for k in list(self.keys()):
            self[k]
        return self

    def values(self):
        return [self[k] for k in self.keys()]

    def items(self):
        return [(k, self[k]) for k in self.keys()]


class Decoder(nn.Module):
    def __init__(self, num_layers, d_model, num_heads, d_ff, dropout, max_len, target_vocab_size,
                 compute_dtype=torch.bfloat16):
        super().__init__()
        if d_model != 768 or d_model // num_heads not in (64, 96):
            raise NotImplementedError("sm_100a decoder kernels are instantiated for d_model 768 with head dim 64 or 96")
        self.max_len = max_len
        self.num_layers = num_layers
        self.num_heads = num_heads
        self.d_model = d_model
        self.dec_layers = nn.ModuleList([DecoderLayer(d_model, num_heads, d_ff, dropout) for _ in range(num_layers)])
        self.target_embedding = nn.Embedding(target_vocab_size, d_model)
        self.positional_encoding = PositionalEncoding(d_model)
        self.final_linear = nn.Linear(d_model, target_vocab_size)
        self.dropout = nn.Dropout(dropout)
        self.compute_dtype = compute_dtype
        self._cache = {}
        self._graphs = {}  # captured greedy programs: key -> (CUDAGraph, static input, static output, kernels per replay)
        self.graph_replayed_launches = 0

    # ------------------------------------------------------------------ teacher-forced forward (fine-tune / val)
    def _layer(self, layer, e2, mem2, B, T, Lm, pad_mask_add, drop_p, seed):
        """DecoderLayer.forward (m3ae_decoder.py:54-90) on [B*T, D] rows.  Train mode applies the reference's dropouts
        (p = `dropout`): on the self- and cross-attention probabilities inside the two nn.MultiheadAttention modules
        (:42-43), after each sub-block before the residual add (dropout1/2/3, :62,:80,:86) — as GEMM-epilogue /
        attention-kernel Philox dropout with per-site streams.  Returns the output rows and what the lazy
        attention-map handles need (q, k of both attentions)."""
        dt = self.compute_dtype
        D, nh = self.d_model, self.num_heads
        dh = D // nh
        xn = Fn.layernorm(e2, layer.pre_norm.weight, layer.pre_norm.bias, layer.pre_norm.eps)
        qkv = Fn.linear(xn, layer.mha1.in_proj_weight, layer.mha1.in_proj_bias, dtype=dt)
        a = Fn.attention_packed(qkv, pad_mask_add, B, nh, T, dh, causal=True, drop_p=drop_p, drop_seed=seed, drop_site=1)
        x = Fn.linear(a, layer.mha1.out_proj.weight, layer.mha1.out_proj.bias, dtype=dt, residual=e2,
                      drop_p=drop_p, drop_seed=seed, drop_site=2)                      # residual + dropout1(mha1)
        residual = x
        x = Fn.layernorm(x, layer.layernorm1.weight, layer.layernorm1.bias, layer.layernorm1.eps)
        E = D
        w2, b2 = layer.mha2.in_proj_weight, layer.mha2.in_proj_bias
        q2 = Fn.linear(x, w2[:E], b2[:E], dtype=dt)
        kv2 = Fn.linear(mem2, w2[E:], b2[E:], dtype=dt)
        c = Fn.attention_packed_kv(q2, kv2, None, B, nh, T, Lm, dh, causal=False, drop_p=drop_p, drop_seed=seed, drop_site=3)
        x = Fn.linear(c, layer.mha2.out_proj.weight, layer.mha2.out_proj.bias, dtype=dt, residual=residual,
                      drop_p=drop_p, drop_seed=seed, drop_site=4)                      # dropout2(mha2) + residual
        residual = x
        x = Fn.layernorm(x, layer.layernorm2.weight, layer.layernorm2.bias, layer.layernorm2.eps)
        x = Fn.ffn(x, layer.ffn[0].weight, layer.ffn[0].bias, layer.ffn[2].weight, layer.ffn[2].bias, residual=residual,
                   act=ops.ACT_RELU, dtype=dt, drop_p=drop_p, drop_seed=seed, drop_site=5)  # dropout3(ffn) + residual
        x = Fn.layernorm(x, layer.layernorm3.weight, layer.layernorm3.bias, layer.layernorm3.eps)
        return x, (qkv.detach()[:, :D], qkv.detach()[:, D:2 * D], q2.detach(), kv2.detach()[:, :D])

    def _avg_probs(self, q, k, B, Lq, Lk, mask, causal):
        """Head-averaged attention weights, as nn.MultiheadAttention(need_weights=True) returns (on demand only)."""
        nh = self.num_heads
        dh = self.d_model // nh
        with torch.no_grad():
            probs = torch.empty(B, nh, Lq, Lk, dtype=torch.float32, device=q.device)
            out = torch.empty(B * Lq, nh * dh, dtype=q.dtype, device=q.device)
            lse = torch.empty(B, nh, Lq, dtype=torch.float32, device=q.device)
            ops.attn_fwd(q, k, k, out, lse, B=B, H=nh, Lq=Lq, Lk=Lk, dh=dh, mask=mask, causal=causal, probs=probs)
            return probs.mean(dim=1)

    def forward(self, padded_targets, padding_mask, cross_attn_feats, return_all_attentions=False):
        """-> (logits [B,T,V] fp32, dict of head-averaged attention maps `layer{i}_dec_self` / `layer{i}_dec_cross` for all
        layers, as the reference returns, m3ae_decoder.py:130-138).  Only dec_layers[-1] reaches the logits (:131-134),
        so only it runs here; the maps are computed when READ (the reference's callers drop them,
        m3ae_decoder.py:358): the live layer's from its q/k, the dead layers' by evaluating that layer then.  Maps are
        the pre-dropout probabilities.  return_all_attentions=True evaluates everything eagerly."""
        if not cross_attn_feats.is_cuda:
            raise RuntimeError("Decoder (B200) needs CUDA tensors: there is no CPU fallback")
        dt = self.compute_dtype
        if torch.is_grad_enabled():
            # a differentiated forward is followed by an optimizer step; fused optimizers do not bump the parameters'
            # version counters, so the greedy path's operand cache cannot be trusted afterwards
            self.invalidate_cache()
        B, T = padded_targets.shape
        Lm, D = cross_attn_feats.shape[1], self.d_model
        dev = cross_attn_feats.device
        pad_mask_add = None
        if padding_mask is not None:  # key_padding_mask = ~padding_mask -> -inf on padded keys (:121)
            pad_mask_add = torch.zeros(B, T, dtype=torch.float32, device=dev).masked_fill(~padding_mask.to(dev), float("-inf"))
        drop_p = float(self.dropout.p) if self.training else 0.0
        seed = Fn.new_seed() if drop_p > 0 else 0
        emb = self.target_embedding(padded_targets)  # [B,T,D] fp32 (autograd through the table)
        e = emb + (emb + self.positional_encoding.pe[:, :T])  # 2*E + pe (:127)
        e2 = Fn.dropout(e.reshape(B * T, D).to(dt), drop_p, seed, 0)  # :128
        mem2 = cross_attn_feats.reshape(B * Lm, D).to(dt)
        live = self.num_layers - 1
        x, (q1, k1, q2, k2) = self._layer(self.dec_layers[live], e2, mem2, B, T, Lm, pad_mask_add, drop_p, seed)
        logits = Fn.linear(x, self.final_linear.weight, self.final_linear.bias, dtype=dt, out_dtype=torch.float32)

        att = _LazyMaps()
        e2d, mem2d = e2.detach(), mem2.detach()
        for i in range(self.num_layers):
            if i == live:
                att.defer(f"layer{i + 1}_dec_self", lambda: self._avg_probs(q1, k1, B, T, T, pad_mask_add, True))
                att.defer(f"layer{i + 1}_dec_cross", lambda: self._avg_probs(q2, k2, B, T, Lm, None, False))
            else:
                def dead(i=i, cache={}):
                    if not cache:
                        with torch.no_grad():
                            _, (a, b, c, d) = self._layer(self.dec_layers[i], e2d, mem2d, B, T, Lm, pad_mask_add, 0.0, 0)
                            cache["s"] = self._avg_probs(a, b, B, T, T, pad_mask_add, True)
                            cache["c"] = self._avg_probs(c, d, B, T, Lm, None, False)
                    return cache
                att.defer(f"layer{i + 1}_dec_self", lambda dead=dead: dead()["s"])
                att.defer(f"layer{i + 1}_dec_cross", lambda dead=dead: dead()["c"])
        if return_all_attentions:
            att.materialize()
        return logits.view(B, T, -1), att

    def invalidate_cache(self):
        """Drop the greedy path's bf16 operand copies and captured CUDA graphs.  Called by every differentiated forward;
        call it yourself after changing parameters in any other way (load_state_dict does it too)."""
        self._cache = {}
        self._graphs = {}

    def load_state_dict(self, *a, **k):
        self.invalidate_cache()
        return super().load_state_dict(*a, **k)

    # ------------------------------------------------------------------ KV-cached greedy decoding
    def _operands(self):
        """Compute-dtype copies of the live layer's weights (refreshed when the parameters change)."""
        layer = self.dec_layers[self.num_layers - 1]
        ps = [layer.mha1.in_proj_weight, layer.mha1.out_proj.weight, layer.mha2.in_proj_weight, layer.mha2.out_proj.weight,
              layer.ffn[0].weight, layer.ffn[2].weight, self.final_linear.weight]
        # (address, version) misses in-place updates by fused optimizers (they do not bump _version): every differentiated
        # forward, load_state_dict and invalidate_cache() clear the cache, and a cheap content digest of one weight row per
        # tensor catches an optimizer step that was applied without any of those (search_path right after opt.step()).
        with torch.no_grad():
            digest = torch.stack([p.detach().reshape(-1)[:: max(1, p.numel() // 64)][:64].double().sum() for p in ps])
            digest = tuple(digest.tolist())
        key = (self.compute_dtype,) + tuple((p.data_ptr(), p._version) for p in ps) + digest
        if self._cache.get("key") != key:
            dt = self.compute_dtype
            with torch.no_grad():
                self._cache = {"key": key, "w": [p.detach() if p.dtype == dt else ops.cast(p.detach().contiguous(), torch.empty_like(p, dtype=dt)) for p in ps]}
        return self._cache["w"]

    @torch.no_grad()
    def search_path(self, cross_attn_feats, tokenizer, early_exit_every=0, use_graph=None):
        """Greedy decoding, same result as the reference's search_path (m3ae_decoder.py:141-182): ids [B, max_len].

        The whole greedy program (memory K/V projection + max_len token steps, ~15 kernels each, every buffer and every
        cache row at a fixed address) is captured ONCE per (shape, operand cache) in a CUDA graph and replayed: issued
        kernel by kernel from Python it is launch-bound (363 us per token step against ~32 us of HBM time, batch 128).
        Parameters read as fp32 masters (biases, LayerNorm, embedding table) are read at replay time; the bf16 operand
        cache is part of the graph key, so retrained weights re-capture.  use_graph=False, M3AE_DECODER_GRAPH=0 or
        early_exit_every > 0 (needs a host sync) run the same program eagerly."""
        if not cross_attn_feats.is_cuda:
            raise RuntimeError("Decoder.search_path (B200) needs CUDA tensors: there is no CPU fallback")
        import os
        if use_graph is None:
            use_graph = os.environ.get("M3AE_DECODER_GRAPH", "1") != "0"
        operands = self._operands()  # outside any capture: checks the masters against the cache (one small D2H read)
        if not use_graph or early_exit_every:
            return self._greedy_program(cross_attn_feats, tokenizer, early_exit_every, operands)
        layer = self.dec_layers[self.num_layers - 1]
        live = list(layer.parameters()) + [self.target_embedding.weight, self.final_linear.bias, self.positional_encoding.pe]
        key = (tuple(cross_attn_feats.shape), cross_attn_feats.dtype, cross_attn_feats.device, self.compute_dtype, self.max_len,
               tokenizer.cls_token_id, tokenizer.sep_token_id, tokenizer.eos_token_id, tokenizer.pad_token_id,
               tuple(w.data_ptr() for w in operands), tuple(p.data_ptr() for p in live))
        entry = self._graphs.get(key)
        if entry is None:
            mem_in = torch.empty_like(cross_attn_feats, memory_format=torch.contiguous_format)
            mem_in.copy_(cross_attn_feats)
            self._greedy_program(mem_in, tokenizer, 0, operands)  # eager warm-up: one-time kernel attribute set-up happens outside capture
            torch.cuda.synchronize(cross_attn_feats.device)
            graph = torch.cuda.CUDAGraph()
            n0 = ops.launch_count()
            with torch.cuda.graph(graph):
                out = self._greedy_program(mem_in, tokenizer, 0, operands)
            n_kernels = ops.launch_count() - n0  # C-ABI kernels recorded in the graph = launched by every replay
            if len(self._graphs) >= 2:  # each graph pins its buffers (cross K/V cache: B*Lm*1536*2 B); keep the two latest
                self._graphs.pop(next(iter(self._graphs)))
            entry = self._graphs[key] = (graph, mem_in, out, n_kernels)
        graph, mem_in, out, n_kernels = entry
        mem_in.copy_(cross_attn_feats)
        graph.replay()
        self.graph_replayed_launches += n_kernels  # ops.launch_count() only sees capture-time calls
        return out.clone()

    @staticmethod
    def _all_finished(nxt, sep, eos):
        """The reference's break condition (m3ae_decoder.py:157-160): every row has produced [SEP] (or eos).  One host sync."""
        done = nxt == sep
        if eos is not None:
            done = done | (nxt == eos)
        return bool(done.any(dim=0).all())

    def _greedy_program(self, cross_attn_feats, tokenizer, early_exit_every=0, operands=None):
        """The greedy decoding program itself (see search_path); max_len single-token steps.  Since the reference blanks
        everything after the first [SEP] and only breaks once ALL rows are finished, running the remaining steps cannot
        change the output (early_exit_every=n adds the reference's all-finished check, one host sync, every n steps)."""
        dt = self.compute_dtype
        dev = cross_attn_feats.device
        B, Lm, D = cross_attn_feats.shape
        nh, L = self.num_heads, self.max_len
        dh = D // nh
        layer = self.dec_layers[self.num_layers - 1]
        w_in1, w_out1, w_in2, w_out2, w_f0, w_f2, w_vocab = operands if operands is not None else self._operands()
        b_in1, b_in2 = layer.mha1.in_proj_bias.detach(), layer.mha2.in_proj_bias.detach()
        V = w_vocab.shape[0]
        sep, pad, eos = tokenizer.sep_token_id, tokenizer.pad_token_id, tokenizer.eos_token_id

        def buf(*shape, dtype=dt):
            return torch.empty(*shape, dtype=dtype, device=dev)

        # cross-attention K/V of the memory: projected once (the reference re-projects them every step)
        mem2 = cross_attn_feats.reshape(B * Lm, D)
        mem2 = mem2 if (mem2.dtype == dt and mem2.is_contiguous()) else ops.cast(mem2.contiguous(), buf(B * Lm, D))
        kv_mem = buf(B, Lm, 2 * D)
        ops.gemm(mem2, w_in2[D:], kv_mem.view(B * Lm, 2 * D), bias=b_in2[D:].contiguous())
        k_mem, v_mem = kv_mem[:, :, :D], kv_mem[:, :, D:]
        k_self, v_self = buf(B, L, D), buf(B, L, D)
        x, xn, q, a, x1, x1n, q2, c, x3, x3n, h, u, x5, x6 = (buf(B, D) for _ in range(14))
        h, u = buf(B, w_f0.shape[0]), buf(B, w_f0.shape[0])
        # the vocabulary projection never writes its [B, V] logits: its epilogue records, per row and 64-column group, the
        # maximum and its column; argmax_partials picks the row argmax (log_softmax is monotone, m3ae_decoder.py:151-153)
        # (bf16 tensor-core path; the fp32 strict-parity path keeps logits + argmax_rows)
        fused_argmax = dt == torch.bfloat16
        n_grp = (V + 63) // 64
        if fused_argmax:
            amax_v, amax_i = buf(n_grp, B, dtype=torch.float32), buf(n_grp, B, dtype=torch.int32)
        else:
            logits = buf(B, V, dtype=torch.float32)
        m1, r1 = buf(B, dtype=torch.float32), buf(B, dtype=torch.float32)
        seq = torch.full((L + 1, B), pad, dtype=torch.int64, device=dev)  # time-major: seq[t] is contiguous
        seq[0] = tokenizer.cls_token_id
        bq, bk, bv = (b_in1[i * D:(i + 1) * D].contiguous() for i in range(3))
        bq2 = b_in2[:D].contiguous()
        pe = self.positional_encoding.pe[0].contiguous()
        table = self.target_embedding.weight.detach()
        pos_all = torch.arange(L, dtype=torch.int32, device=dev).repeat_interleave(B).view(L, B)
        # bf16, batch <= 128: the six per-token projections run on the skinny kernel (whole activation panel + weight slice in
        # shared memory, LayerNorm folded into the load, K/V written straight into the cache rows): csrc/skinny_linear.cu
        import os
        skinny = dt == torch.bfloat16 and B <= 128 and os.environ.get("M3AE_DECODER_SKINNY", "1") != "0"
        if skinny:
            f32 = torch.float32
            fp = lambda v: v.detach().to(f32).contiguous()  # noqa: E731
            ws = torch.zeros(128 * D, dtype=f32, device=dev)
            ctr = torch.zeros(256, dtype=torch.int32, device=dev)
            # LayerNorm folded into the projection's activation load only for small batches: every CTA normalises all B rows
            # redundantly (measured at B = 128: +7 us per launch against 4.3 us for the separate LayerNorm launch)
            fuse_ln = B <= 32
            norms = [(layer.pre_norm, x, xn), (layer.layernorm1, x1, x1n), (layer.layernorm2, x3, x3n)]
            ln_arg = [(fp(n.weight), fp(n.bias), n.eps) if fuse_ln else None for n, _, _ in norms]
            src = [raw if fuse_ln else normed for _, raw, normed in norms]
            S_qkv = ops.SkinnyLinear(src[0], w_in1, [q, k_self[:, 0, :], v_self[:, 0, :]], bias=fp(b_in1), ln=ln_arg[0])
            S_o1 = ops.SkinnyLinear(a, w_out1, x1, bias=fp(layer.mha1.out_proj.bias), residual=x)
            S_q2 = ops.SkinnyLinear(src[1], w_in2[:D], q2, bias=bq2.to(f32).contiguous(), ln=ln_arg[1])
            S_o2 = ops.SkinnyLinear(c, w_out2, x3, bias=fp(layer.mha2.out_proj.bias), residual=x1)
            S_f0 = ops.SkinnyLinear(src[2], w_f0, h, bias=fp(layer.ffn[0].bias), act=ops.ACT_RELU, ln=ln_arg[2])
            S_f2 = ops.SkinnyLinear(h, w_f2, x5, bias=fp(layer.ffn[2].bias), residual=x3, workspace=ws, counters=ctr)
            k0_ptr, v0_ptr, row_bytes = k_self.data_ptr(), v_self.data_ptr(), D * k_self.element_size()

            def norm(i):
                if not fuse_ln:
                    n, raw, normed = norms[i]
                    ops.layernorm_fwd(raw, n.weight, n.bias, n.eps, normed, m1, r1)
        if skinny:
            pn = (fp(layer.pre_norm.weight), fp(layer.pre_norm.bias), layer.pre_norm.eps)
            # token 0: embedding of [CLS] + pre_norm; later tokens: argmax + embedding + pre_norm in one launch (below)
            ops.decoder_next_token(x, None if fuse_ln else xn, table, pe, 0, ids_in=seq[0], ln=pn)
        for t in range(L):
            if skinny:
                S_qkv.p.out[1], S_qkv.p.out[2] = k0_ptr + t * row_bytes, v0_ptr + t * row_bytes  # cache rows [b, t, :]
                S_qkv.run()
                ops.decode_attn(q, k_self, v_self, a, H=nh, length=t + 1, dh=dh)
                S_o1.run()
                norm(1)
                S_q2.run()
                ops.decode_attn(q2, k_mem, v_mem, c, H=nh, length=Lm, dh=dh)
                S_o2.run()
                norm(2)
                S_f0.run()
                S_f2.run()
                ops.layernorm_fwd(x5, layer.layernorm3.weight, layer.layernorm3.bias, layer.layernorm3.eps, x6, m1, r1)
                ops.gemm_group([ops.GemmSpec(x6, w_vocab, None, bias=self.final_linear.bias, amax=(amax_v, amax_i))])
                if t + 1 < L:
                    ops.decoder_next_token(x, None if fuse_ln else xn, table, pe, t + 1, ids_out=seq[t + 1], amax=(amax_v, amax_i), ln=pn)
                else:
                    ops.argmax_partials(amax_v, amax_i, seq[t + 1])
                if early_exit_every and (t + 1) % early_exit_every == 0 and self._all_finished(seq[1:t + 2], sep, eos):
                    break
                continue
            ops.decoder_embed(seq[t], pos_all[t], table, pe, x)  # 2*E[id] + pe[t]
            ops.layernorm_fwd(x, layer.pre_norm.weight, layer.pre_norm.bias, layer.pre_norm.eps, xn, m1, r1)
            # q for this token; k, v written straight into the cache rows [b, t, :]
            ops.gemm_group([ops.GemmSpec(xn, w_in1[:D], q, bias=bq),
                            ops.GemmSpec(xn, w_in1[D:2 * D], k_self[:, t, :], bias=bk),
                            ops.GemmSpec(xn, w_in1[2 * D:], v_self[:, t, :], bias=bv)])
            ops.decode_attn(q, k_self, v_self, a, H=nh, length=t + 1, dh=dh)
            ops.gemm(a, w_out1, x1, bias=layer.mha1.out_proj.bias, residual=x)
            ops.layernorm_fwd(x1, layer.layernorm1.weight, layer.layernorm1.bias, layer.layernorm1.eps, x1n, m1, r1)
            ops.gemm(x1n, w_in2[:D], q2, bias=bq2)
            ops.decode_attn(q2, k_mem, v_mem, c, H=nh, length=Lm, dh=dh)
            ops.gemm(c, w_out2, x3, bias=layer.mha2.out_proj.bias, residual=x1)
            ops.layernorm_fwd(x3, layer.layernorm2.weight, layer.layernorm2.bias, layer.layernorm2.eps, x3n, m1, r1)
            ops.gemm(x3n, w_f0, h, bias=layer.ffn[0].bias, act=ops.ACT_RELU)
            ops.gemm(h, w_f2, x5, bias=layer.ffn[2].bias, residual=x3)
            ops.layernorm_fwd(x5, layer.layernorm3.weight, layer.layernorm3.bias, layer.layernorm3.eps, x6, m1, r1)
            if fused_argmax:
                ops.gemm_group([ops.GemmSpec(x6, w_vocab, None, bias=self.final_linear.bias, amax=(amax_v, amax_i))])
                ops.argmax_partials(amax_v, amax_i, seq[t + 1])
            else:
                ops.gemm(x6, w_vocab, logits, bias=self.final_linear.bias)
                ops.argmax_rows(logits, seq[t + 1])
            if early_exit_every and (t + 1) % early_exit_every == 0 and self._all_finished(seq[1:t + 2], sep, eos):
                break
        out = seq[1:].t().contiguous()
        special = out == sep
        if eos is not None:
            special = special | (out == eos)
        after = (special.cumsum(dim=1) - special.long()) > 0  # strictly after the first [SEP]/[EOS] (:165-173)
        return out.masked_fill(after, pad)
