"""ORACLE — test infrastructure only (never imported by the product path).

CPU fp32 restatement of the reference's hot path (better62/MM-VQA-Healthcare), written with plain torch
CPU tensor ops (matmul / softmax / layer_norm — the same library calls the reference's own CPU path makes),
one function per reference function, each citing the file:line it follows (paths relative to the
reference root).  Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline legs may import it.

Pinning: the reference ships no tests, golden vectors or fixtures, so this restatement is
pinned against OUTPUTS OF THE REFERENCE ITSELF: `oracle/make_golden.py` imports the unmodified reference
modules from /root/reference (via `oracle/ref_shim.py`), runs them on seeded inputs/weights produced by the
helpers at the bottom of this file and stores the results under `tests/golden/`;
`tests/test_oracle_golden.py` checks this file against those vectors.

Weights are passed as a flat `state_dict`-style mapping using the reference's own key names.
"""
import math

import torch
import torch.nn.functional as F

# ----------------------------------------------------------------------------- BertCrossLayer subtree


def linear(x, w, b=None):
    """nn.Linear: y = x W^T + b (weight layout [out, in])."""
    y = x @ w.t()
    return y if b is None else y + b


def extended_attention_mask(mask):
    """HF 4.6.0 ModuleUtilsMixin.get_extended_attention_mask as called at m3ae_module.py:232,255:
    [B,L] {0,1} -> additive [B,1,1,L] with (1 - mask) * -10000.0."""
    return (1.0 - mask[:, None, None, :].to(torch.float32)) * -10000.0


def bert_self_attention(hidden, sd, prefix, num_heads, attention_mask=None, encoder_hidden=None,
                        encoder_attention_mask=None, dropout_p=0.0, training=False):
    """BertSelfAttention.forward, bert_model.py:253-350 (self mode :284-286, cross mode :275-278)."""
    B, Lq, Hd = hidden.shape
    dh = Hd // num_heads
    q = linear(hidden, sd[prefix + "query.weight"], sd[prefix + "query.bias"])  # :263
    kv_src = hidden if encoder_hidden is None else encoder_hidden
    mask = attention_mask if encoder_hidden is None else encoder_attention_mask  # :278
    k = linear(kv_src, sd[prefix + "key.weight"], sd[prefix + "key.bias"])
    v = linear(kv_src, sd[prefix + "value.weight"], sd[prefix + "value.bias"])

    def split(x):  # transpose_for_scores :248-251
        return x.view(x.shape[0], x.shape[1], num_heads, dh).permute(0, 2, 1, 3)

    q, k, v = split(q), split(k), split(v)
    scores = q @ k.transpose(-1, -2)  # :301
    scores = scores / math.sqrt(dh)  # :319
    if mask is not None:
        scores = scores + mask  # :322
    probs = torch.softmax(scores, dim=-1)  # :325
    probs = F.dropout(probs, dropout_p, training)  # :334
    ctx = probs @ v  # :340
    ctx = ctx.permute(0, 2, 1, 3).contiguous().view(B, Lq, Hd)  # :342-344
    return ctx, probs


def bert_self_output(ctx, residual, sd, prefix, eps, dropout_p=0.0, training=False):
    """BertSelfOutput.forward :360-364 and BertOutput.forward :438-442 (same arithmetic)."""
    h = linear(ctx, sd[prefix + "dense.weight"], sd[prefix + "dense.bias"])
    h = F.dropout(h, dropout_p, training)
    return F.layer_norm(h + residual, (h.shape[-1],), sd[prefix + "LayerNorm.weight"], sd[prefix + "LayerNorm.bias"], eps)


def bert_attention(hidden, sd, prefix, num_heads, eps, attention_mask=None, encoder_hidden=None,
                   encoder_attention_mask=None, dropout_p=0.0, training=False):
    """BertAttention.forward :392-413."""
    ctx, probs = bert_self_attention(hidden, sd, prefix + "self.", num_heads, attention_mask, encoder_hidden,
                                     encoder_attention_mask, dropout_p, training)
    return bert_self_output(ctx, hidden, sd, prefix + "output.", eps, dropout_p, training), probs


def bert_intermediate(x, sd, prefix):
    """BertIntermediate.forward :425-428 — exact erf GELU (ACT2FN["gelu"], :421)."""
    return F.gelu(linear(x, sd[prefix + "dense.weight"], sd[prefix + "dense.bias"]))


def bert_cross_layer(hidden, encoder_hidden, sd, prefix="", attention_mask=None, encoder_attention_mask=None,
                     num_heads=12, eps=1e-12, dropout_p=0.0, training=False):
    """BertCrossLayer.forward + feed_forward_chunk, bert_model.py:457-503.
    Returns (layer_output, self_probs, cross_probs) — the output_attentions=True tuple."""
    a1, self_probs = bert_attention(hidden, sd, prefix + "attention.", num_heads, eps, attention_mask,
                                    dropout_p=dropout_p, training=training)  # :467-473
    a2, cross_probs = bert_attention(a1, sd, prefix + "crossattention.", num_heads, eps, attention_mask,
                                     encoder_hidden, encoder_attention_mask, dropout_p, training)  # :480-488
    inter = bert_intermediate(a2, sd, prefix + "intermediate.")  # :499
    out = bert_self_output(inter, a2, sd, prefix + "output.", eps, dropout_p, training)  # :500
    return out, self_probs, cross_probs


# ----------------------------------------------------------------------------- M3AETransformerSS.infer fusion section


def pooler(x, sd, prefix):
    """prediction_heads.Pooler.forward, prediction_heads.py:15-19."""
    return torch.tanh(linear(x[:, 0], sd[prefix + "dense.weight"], sd[prefix + "dense.bias"]))


def fusion_forward(uni_text_feats, uni_image_feats, text_masks, sd, num_layers=6, num_heads=12, eps=1e-12,
                   image_token_type_idx=1, dropout_p=0.0, training=False, output_attentions=False):
    """Fusion section of M3AETransformerSS.infer, m3ae_module.py:235,252-297.
    Inputs are the uni-modal encoder outputs (RoBERTa last layer [B,Lt,768]; CLIP ViT tokens [B,Li,768])."""
    x = linear(uni_text_feats, sd["multi_modal_language_proj.weight"], sd["multi_modal_language_proj.bias"])  # :235
    y = linear(uni_image_feats, sd["multi_modal_vision_proj.weight"], sd["multi_modal_vision_proj.bias"])  # :252
    ext_text = extended_attention_mask(text_masks)  # :232
    image_masks = torch.ones(y.shape[0], y.shape[1], dtype=torch.long)  # :253-254
    ext_image = extended_attention_mask(image_masks)  # :255
    tte = sd["modality_type_embeddings.weight"]
    x = x + tte[0]  # :261
    y = y + tte[image_token_type_idx]  # :262
    attns = {"text2image_attns": [], "image2text_attns": []}
    for i in range(num_layers):  # :268-285 — both layers read the OLD x, y
        x1 = bert_cross_layer(x, y, sd, f"multi_modal_language_layers.{i}.", ext_text, ext_image, num_heads, eps,
                              dropout_p, training)
        y1 = bert_cross_layer(y, x, sd, f"multi_modal_vision_layers.{i}.", ext_image, ext_text, num_heads, eps,
                              dropout_p, training)
        x, y = x1[0], y1[0]
        if output_attentions:
            attns["text2image_attns"].append(x1[1:])
            attns["image2text_attns"].append(y1[1:])
    cls_text = pooler(x, sd, "multi_modal_language_pooler.")  # :289
    cls_image = pooler(y, sd, "multi_modal_vision_pooler.")  # :291
    cls = torch.cat([cls_text, cls_image], dim=-1)  # :296
    return {
        "extended_text_masks": ext_text,
        "extended_image_masks": ext_image,
        "multi_modal_text_feats": x,
        "multi_modal_image_feats": y,
        "multi_modal_cls_feats": cls,
        "attentions": attns if output_attentions else None,
    }


def vqa_head(cls_feats, sd, prefix="vqa_head."):
    """nn.Sequential(Linear, LayerNorm(eps 1e-5), GELU, Linear), m3ae_module.py:120-125."""
    h = linear(cls_feats, sd[prefix + "0.weight"], sd[prefix + "0.bias"])
    h = F.layer_norm(h, (h.shape[-1],), sd[prefix + "1.weight"], sd[prefix + "1.bias"], 1e-5)
    h = F.gelu(h)
    return linear(h, sd[prefix + "3.weight"], sd[prefix + "3.bias"])


def vqa_targets(vqa_labels, vqa_scores, num_labels):
    """Multi-hot target construction, objectives.py:188,195-197."""
    t = torch.zeros(len(vqa_labels), num_labels)
    for i, (labels, scores) in enumerate(zip(vqa_labels, vqa_scores)):
        for l, s in zip(labels, scores):
            t[i, l] = s
    return t


def vqa_loss(logits, targets):
    """objectives.py:201: BCE-with-logits (mean) * number of labels."""
    return F.binary_cross_entropy_with_logits(logits, targets) * targets.shape[1]


def finetune_step(uni_text_feats, uni_image_feats, text_masks, targets, sd, dropout_p=0.0, training=False):
    """compute_vqa_m3ae restricted to the hot path (objectives.py:176-201): fusion -> vqa_head -> loss."""
    ret = fusion_forward(uni_text_feats, uni_image_feats, text_masks, sd, dropout_p=dropout_p, training=training)
    logits = vqa_head(ret["multi_modal_cls_feats"], sd)
    return vqa_loss(logits, targets), logits, ret


# ----------------------------------------------------------------------------- generative decoder head


def causal_mask(T):
    """CausalMask, m3ae_decoder.py:14-20: True above the diagonal (= not allowed)."""
    return ~torch.tril(torch.ones((T, T), dtype=torch.bool), diagonal=0)


def positional_encoding_table(d_model, max_len=1024):
    """PositionalEncoding.__init__, m3ae_decoder.py:24-34 -> pe [1,max_len,d_model]."""
    pe = torch.zeros(max_len, d_model)
    position = torch.arange(0, max_len, dtype=torch.float).unsqueeze(1)
    div_term = torch.exp(torch.arange(0, d_model, 2).float() * (-math.log(10000.0) / d_model))
    pe[:, 0::2] = torch.sin(position * div_term)
    pe[:, 1::2] = torch.cos(position * div_term)
    return pe.unsqueeze(0)


def multihead_attention(query, key, value, sd, prefix, num_heads, attn_mask=None, key_padding_mask=None):
    """nn.MultiheadAttention(batch_first=True) forward with need_weights=True (default), eval mode.
    Packed in_proj_weight [3E,E]; scores scaled by 1/sqrt(dh); bool masks fill -inf; returns head-averaged probs.
    (torch/nn/functional.py multi_head_attention_forward; called at m3ae_decoder.py:60-61,78-79.)"""
    B, Lq, E = query.shape
    Lk = key.shape[1]
    dh = E // num_heads
    w, b = sd[prefix + "in_proj_weight"], sd[prefix + "in_proj_bias"]
    q = linear(query, w[:E], b[:E])
    k = linear(key, w[E:2 * E], b[E:2 * E])
    v = linear(value, w[2 * E:], b[2 * E:])
    q = q.view(B, Lq, num_heads, dh).transpose(1, 2)
    k = k.view(B, Lk, num_heads, dh).transpose(1, 2)
    v = v.view(B, Lk, num_heads, dh).transpose(1, 2)
    scores = (q * (1.0 / math.sqrt(dh))) @ k.transpose(-1, -2)
    if attn_mask is not None:
        scores = scores.masked_fill(attn_mask[None, None], float("-inf"))
    if key_padding_mask is not None:
        scores = scores.masked_fill(key_padding_mask[:, None, None, :], float("-inf"))
    probs = torch.softmax(scores, dim=-1)
    ctx = (probs @ v).transpose(1, 2).reshape(B, Lq, E)
    out = linear(ctx, sd[prefix + "out_proj.weight"], sd[prefix + "out_proj.bias"])
    return out, probs.mean(dim=1)


def _ln(x, sd, prefix, eps=1e-5):
    return F.layer_norm(x, (x.shape[-1],), sd[prefix + "weight"], sd[prefix + "bias"], eps)


def decoder_layer(x, memory, sd, prefix, num_heads, pad_mask_dec, slf_attn_mask):
    """DecoderLayer.forward, m3ae_decoder.py:54-90 (eval mode: dropouts are identity).
    Mixed pre/post-norm: the residual is taken BEFORE layernorm1/layernorm2."""
    residual = x
    x_norm = _ln(x, sd, prefix + "pre_norm.")  # :58
    a, w1 = multihead_attention(x_norm, x_norm, x_norm, sd, prefix + "mha1.", num_heads, slf_attn_mask, pad_mask_dec)  # :60
    x = residual + a  # :62
    residual = x
    x = _ln(x, sd, prefix + "layernorm1.")  # :64
    enc_pad = torch.zeros(memory.shape[0], memory.shape[1], dtype=torch.bool)  # :72-76
    c, w2 = multihead_attention(x, memory, memory, sd, prefix + "mha2.", num_heads, None, enc_pad)  # :78
    x = c + residual  # :80-81
    residual = x
    x = _ln(x, sd, prefix + "layernorm2.")  # :83
    f = linear(torch.relu(linear(x, sd[prefix + "ffn.0.weight"], sd[prefix + "ffn.0.bias"])),
               sd[prefix + "ffn.2.weight"], sd[prefix + "ffn.2.bias"])  # :85
    x = f + residual  # :86-87
    x = _ln(x, sd, prefix + "layernorm3.")  # :88
    return x, w1, w2


def decoder_forward(padded_targets, padding_mask, memory, sd, num_layers=6, num_heads=8, all_layers=False):
    """Decoder.forward, m3ae_decoder.py:119-138.
    `target_embed += positional_encoding(target_embed)` gives 2*E[ids] + pe (:127 with :36), and EVERY layer is fed
    target_embed while only the last iteration's x survives (:131-134): output == final_linear(dec_layers[-1](e, mem)).
    all_layers=True also evaluates the dead layers 0..n-2 exactly as the reference does (for the timed CPU baseline
    and for the attention-weights dict)."""
    pad_mask_dec = ~padding_mask if padding_mask is not None else None  # :121
    T = padded_targets.shape[1]
    cm = causal_mask(T)  # :123
    e = sd["target_embedding.weight"][padded_targets]  # :125
    e = e + (e + sd["positional_encoding.pe"][:, :T])  # :127
    att = {}
    x = None
    layers = range(num_layers) if all_layers else [num_layers - 1]
    for i in layers:  # :131-134
        x, att[f"layer{i + 1}_dec_self"], att[f"layer{i + 1}_dec_cross"] = decoder_layer(
            e, memory, sd, f"dec_layers.{i}.", num_heads, pad_mask_dec, cm)
    return linear(x, sd["final_linear.weight"], sd["final_linear.bias"]), att  # :136


def search_path(memory, sd, max_len, cls_id=101, sep_id=102, eos_id=None, pad_id=0, num_layers=6, num_heads=8,
                all_layers=False):
    """Decoder.search_path, m3ae_decoder.py:141-182 (greedy; re-runs the whole prefix each step; stops only when
    ALL rows are finished; blanks everything after the first [SEP]; right-pads to max_len)."""
    B = memory.shape[0]
    seq = torch.full((B, 1), cls_id, dtype=torch.long)
    finished = torch.zeros(B, dtype=torch.bool)
    for _ in range(max_len):  # :148
        logits, _ = decoder_forward(seq, None, memory, sd, num_layers, num_heads, all_layers)
        nxt = torch.log_softmax(logits[:, -1], dim=1).argmax(dim=-1)  # :151-153
        fin = nxt == sep_id
        if eos_id is not None:
            fin = fin | (nxt == eos_id)
        finished |= fin  # :156-158
        seq = torch.cat([seq, nxt[:, None]], dim=-1)
        if finished.all():
            break
    seq = seq[:, 1:]  # :163
    for i in range(B):  # :165-173
        special = seq[i] == sep_id
        if eos_id is not None:
            special = special | (seq[i] == eos_id)
        idx = torch.where(special)[0]
        if len(idx) > 0:
            seq[i, idx[0] + 1:] = pad_id
    return F.pad(seq, (0, max_len - seq.shape[1]), value=pad_id)  # :175-180


# ----------------------------------------------------------------------------- seeded weights / inputs (shared by golden + tests + bench)

CROSS_LAYER_KEYS = [
    ("attention.self.query", 768, 768), ("attention.self.key", 768, 768), ("attention.self.value", 768, 768),
    ("attention.output.dense", 768, 768),
    ("crossattention.self.query", 768, 768), ("crossattention.self.key", 768, 768), ("crossattention.self.value", 768, 768),
    ("crossattention.output.dense", 768, 768),
    ("intermediate.dense", 3072, 768), ("output.dense", 768, 3072),
]
CROSS_LAYER_NORMS = ["attention.output.LayerNorm", "crossattention.output.LayerNorm", "output.LayerNorm"]


def _lin(sd, g, name, out_f, in_f, wstd=0.02, bstd=0.02):
    sd[name + ".weight"] = torch.randn(out_f, in_f, generator=g) * wstd
    sd[name + ".bias"] = torch.randn(out_f, generator=g) * bstd


def _norm(sd, g, name, n, wkey="weight", bkey="bias"):
    sd[f"{name}.{wkey}"] = 1.0 + 0.1 * torch.randn(n, generator=g)
    sd[f"{name}.{bkey}"] = 0.1 * torch.randn(n, generator=g)


def make_cross_layer_state(seed, prefix="", g=None):
    """Seeded BertCrossLayer state_dict (26 tensors, key names of bert_model.py:445-455).  Weights N(0,0.02) as
    init_weights (m3ae_utils.py:101-109); biases and LayerNorm affine are perturbed (not 0/1) so parity tests see them."""
    g = g or torch.Generator().manual_seed(seed)
    sd = {}
    for name, o, i in CROSS_LAYER_KEYS:
        _lin(sd, g, prefix + name, o, i)
    for name in CROSS_LAYER_NORMS:
        _norm(sd, g, prefix + name, 768)
    return sd


def make_fusion_state(seed, num_layers=6, num_labels=498):
    """Seeded state_dict of everything on the fusion path of M3AETransformerSS (m3ae_module.py:70-88,117-125)."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    _lin(sd, g, "multi_modal_language_proj", 768, 768)
    _lin(sd, g, "multi_modal_vision_proj", 768, 768)
    sd["modality_type_embeddings.weight"] = torch.randn(2, 768, generator=g) * 0.02
    for i in range(num_layers):
        sd.update(make_cross_layer_state(0, f"multi_modal_vision_layers.{i}.", g))
    for i in range(num_layers):
        sd.update(make_cross_layer_state(0, f"multi_modal_language_layers.{i}.", g))
    _lin(sd, g, "multi_modal_vision_pooler.dense", 768, 768)
    _lin(sd, g, "multi_modal_language_pooler.dense", 768, 768)
    _lin(sd, g, "vqa_head.0", 1536, 1536)
    _norm(sd, g, "vqa_head.1", 1536)
    _lin(sd, g, "vqa_head.3", num_labels, 1536)
    return sd


def make_decoder_state(seed, num_layers=6, d_model=768, d_ff=3072, vocab=30522, live_only=False):
    """Seeded Decoder state_dict (m3ae_decoder.py:94-116).  live_only skips dec_layers.0..n-2 (dead compute)."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for i in range(num_layers):
        p = f"dec_layers.{i}."
        layer = {}
        for m in ("mha1", "mha2"):
            layer[p + m + ".in_proj_weight"] = torch.randn(3 * d_model, d_model, generator=g) * 0.03
            layer[p + m + ".in_proj_bias"] = torch.randn(3 * d_model, generator=g) * 0.02
            _lin(layer, g, p + m + ".out_proj", d_model, d_model, 0.03)
        _lin(layer, g, p + "ffn.0", d_ff, d_model, 0.03)
        _lin(layer, g, p + "ffn.2", d_model, d_ff, 0.03)
        for n in ("pre_norm", "layernorm1", "layernorm2", "layernorm3"):
            _norm(layer, g, p + n, d_model)
        if not live_only or i == num_layers - 1:
            sd.update(layer)
    sd["target_embedding.weight"] = torch.randn(vocab, d_model, generator=g)
    sd["positional_encoding.pe"] = positional_encoding_table(d_model)
    _lin(sd, g, "final_linear", vocab, d_model, 0.05)
    return sd


def seeded_tensor_like(name, t, seed):
    """Deterministic stand-in weights for a named state-dict tensor: the same (name, shape, seed) gives the same values in
    the reference model and in the drop-in, so whole-model parity fixtures need not ship 1.3 GB of weights.
    >= 2-D tensors ~ N(0, 0.03); LayerNorm-style `weight` vectors ~ 1 + 0.1 N(0,1); other vectors ~ 0.02 N(0,1)."""
    import zlib
    g = torch.Generator().manual_seed((zlib.crc32(name.encode()) + 7919 * seed) & 0x7FFFFFFF)
    if t.dim() >= 2:
        return torch.randn(t.shape, generator=g) * 0.03
    if name.endswith("weight"):
        return 1.0 + 0.1 * torch.randn(t.shape, generator=g)
    return 0.02 * torch.randn(t.shape, generator=g)


def seed_model_parameters(model, seed):
    """In-place: every floating-point PARAMETER of `model` := seeded_tensor_like(its state-dict name)."""
    with torch.no_grad():
        for name, p in model.named_parameters():
            if p.dtype.is_floating_point:
                p.copy_(seeded_tensor_like(name, p, seed))
    return model


def make_model_batch(seed, B, image_size=288, Lt=32, num_labels=498, device="cpu"):
    """Seeded synthetic batch dict in the layout BaseDataset.collate produces (base_dataset.py:165-228)."""
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(3, 50265, (B, Lt), generator=g)
    ids[:, 0] = 0
    lens = torch.randint(8, Lt + 1, (B,), generator=g)
    masks = (torch.arange(Lt)[None, :] < lens[:, None]).long()
    ids = torch.where(masks.bool(), ids, torch.ones_like(ids))
    labels = torch.randint(0, num_labels, (B,), generator=g)
    image = torch.randn(B, 3, image_size, image_size, generator=g)
    return {"image": [image.to(device)], "text_ids": ids.to(device), "text_masks": masks.to(device),
            "text_labels": torch.full((B, Lt), -100).to(device), "vqa_labels": [[int(l)] for l in labels],
            "vqa_scores": [[1.0]] * B, "answer_types": [int(l) % 2 for l in labels],
            "vqa_answer": [[f"answer {int(l)}"] for l in labels]}


class SynthTokenizer:
    """Stand-in for HF BertTokenizer('bert-base-uncased') (m3ae_decoder.py:191; no vocabulary file offline): whitespace
    split, hashed ids in [1000, vocab), BERT's special ids.  Same surface the reference uses (:336-342, :326-329)."""

    cls_token_id, sep_token_id, pad_token_id, eos_token_id = 101, 102, 0, None

    def __init__(self, vocab_size=30522):
        self.vocab_size = vocab_size
        self._words = {}

    def _id(self, w):
        h = 1000 + (sum((i + 1) * ord(c) * 2654435761 for i, c in enumerate(w)) % (self.vocab_size - 1000))
        self._words.setdefault(h, w)
        return h

    def __call__(self, texts, padding=True, truncation=True, return_tensors="pt", max_length=12):
        rows = [([self.cls_token_id] + [self._id(w) for w in t.lower().split()])[:max_length - 1] + [self.sep_token_id] for t in texts]
        width = max(len(r) for r in rows)
        out = torch.full((len(rows), width), self.pad_token_id, dtype=torch.long)
        for i, r in enumerate(rows):
            out[i, :len(r)] = torch.tensor(r)
        return type("Enc", (), {"input_ids": out})()

    def decode(self, seq, skip_special_tokens=True):
        ids = seq.tolist() if torch.is_tensor(seq) else list(seq)
        special = {self.cls_token_id, self.sep_token_id, self.pad_token_id}
        return " ".join(self._words.get(i, f"[{i}]") for i in ids if not (skip_special_tokens and i in special))


class SynthT5Tokenizer:
    """Stand-in for HF T5Tokenizer('t5-small') (m3ae_t5_mm_encoder_input.py:26; no sentencepiece model offline): whitespace
    split, hashed ids in [3, 32000), T5's special ids (pad 0, eos 1).  Same surface the reference uses (:113-117, :232-237,
    :262-265): `__call__(str | list[str], return_tensors, add_special_tokens, padding, truncation).input_ids`, `batch_decode`."""

    pad_token_id, eos_token_id, unk_token_id = 0, 1, 2

    def __init__(self, vocab_size=32128):
        self.vocab_size = vocab_size
        self._words = {}

    def _id(self, w):
        h = 3 + (sum((i + 1) * ord(c) * 2654435761 for i, c in enumerate(w)) % 31997)
        self._words.setdefault(h, w)
        return h

    def __call__(self, texts, return_tensors="pt", add_special_tokens=True, padding=False, truncation=False, max_length=512):
        if isinstance(texts, str):
            texts = [texts]
        rows = []
        for t in texts:
            ids = [self._id(w) for w in t.lower().split()]
            if add_special_tokens:
                ids = ids[:max_length - 1] + [self.eos_token_id]
            rows.append(ids)
        width = max(len(r) for r in rows)
        out = torch.full((len(rows), width), self.pad_token_id, dtype=torch.long)
        for i, r in enumerate(rows):
            out[i, :len(r)] = torch.tensor(r, dtype=torch.long)
        return type("Enc", (), {"input_ids": out})()

    def batch_decode(self, seqs, skip_special_tokens=True, **kw):
        outs = []
        for seq in seqs:
            ids = seq.tolist() if torch.is_tensor(seq) else list(seq)
            outs.append(" ".join(self._words.get(i, f"<{i}>") for i in ids if not (skip_special_tokens and i in (0, 1, 2))))
        return outs


def make_fusion_inputs(seed, B, Lt=32, Li=325, num_labels=498, all_valid=False):
    """Seeded synthetic hot-path inputs: uni-modal features of the shapes RoBERTa / CLIP ViT-B/16
    produce at 288 px / 32 tokens, ragged text masks (valid length uniform in [8,32]), one-hot VQA targets."""
    g = torch.Generator().manual_seed(seed)
    text = torch.randn(B, Lt, 768, generator=g)
    image = torch.randn(B, Li, 768, generator=g)
    lens = torch.full((B,), Lt) if all_valid else torch.randint(8, Lt + 1, (B,), generator=g)
    masks = (torch.arange(Lt)[None, :] < lens[:, None]).long()
    labels = torch.randint(0, num_labels, (B,), generator=g)
    targets = vqa_targets([[int(l)] for l in labels], [[1.0]] * B, num_labels)
    return {"text_feats": text, "image_feats": image, "text_masks": masks, "labels": labels, "targets": targets}
