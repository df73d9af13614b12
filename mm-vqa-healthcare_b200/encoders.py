"""Uni-modal encoders of M3AE on the sm_100a kernels.

`ClipVisionEncoder`  the reference's `vision_encoder` = CLIP whose forward is `visual(image)`
                     (m3ae/modules/vision_encoders/clip_model.py:79-128,186-187): 16x16 patch-embed conv as ONE GEMM,
                     class token + positional embedding, ln_pre, 11 pre-LN ResidualAttentionBlocks (packed in-proj MHA,
                     QuickGELU MLP, clip_model.py:41-63,71), ln_post.  Parameter names = the reference's
                     (`visual.conv1.weight`, `visual.transformer.resblocks.{i}.attn.in_proj_weight`, ...), so reference
                     checkpoints load.  LayerNorm statistics are fp32 (the reference's LayerNorm subclass casts to fp32,
                     clip_model.py:27-33); the residual stream is kept in the compute dtype.
`roberta_layer`      one HF `RobertaLayer` (post-LN BERT layer; called one by one with the additive mask at
                     m3ae_module.py:230-234) driven from the HF module's own Parameters: fused QKV GEMM, flash
                     attention with the key-padding mask, out-proj + dropout + residual epilogue, LayerNorm, GELU FFN —
                     as ONE fused autograd node per layer (functional.cross_layers_apply with LayerConfig.has_cross=False:
                     a RobertaLayer is a BertCrossLayer without its cross-attention block).

The ViT's residual attention blocks run as ONE fused autograd node each (vit_block.py; the patch embedding, ln_pre and ln_post
are composable nodes of `functional`); every matmul, attention, LayerNorm and activation on both paths is a kernel of
libm3ae_b200.so; torch is used for the reshapes that build the patch matrix and for the embedding gathers.
"""
import os
from collections import OrderedDict

import torch
import torch.nn as nn

from . import functional as Fn
from . import ops
from .vit_block import vit_block

# M3AE_VIT_FUSED=0: run the CLIP blocks from the composable autograd nodes (A/B and cross-check of the fused node)
VIT_FUSED = os.environ.get("M3AE_VIT_FUSED", "1") != "0"


# ----------------------------------------------------------------------------- CLIP ViT-B/16 visual tower
class _ResBlock(nn.Module):
    """Parameter container of clip_model.ResidualAttentionBlock (:41-52); math in ClipVisionEncoder.forward."""

    def __init__(self, d_model, n_head):
        super().__init__()
        self.attn = nn.MultiheadAttention(d_model, n_head)
        self.ln_1 = nn.LayerNorm(d_model)
        self.mlp = nn.Sequential(OrderedDict([("c_fc", nn.Linear(d_model, d_model * 4)), ("gelu", nn.Identity()),
                                              ("c_proj", nn.Linear(d_model * 4, d_model))]))
        self.ln_2 = nn.LayerNorm(d_model)
        self.n_head = n_head


class _Transformer(nn.Module):
    def __init__(self, width, layers, heads):
        super().__init__()
        self.width, self.layers = width, layers
        self.resblocks = nn.Sequential(*[_ResBlock(width, heads) for _ in range(layers - 1)])  # clip_model.py:71: layers-1


class _VisualTransformer(nn.Module):
    """Parameter container of clip_model.VisualTransformer (:79-93)."""

    def __init__(self, patch_size, width, layers, heads, resolution_after):
        super().__init__()
        self.patch_size = patch_size
        self.conv1 = nn.Conv2d(3, width, kernel_size=patch_size, stride=patch_size, bias=False)
        scale = width ** -0.5
        self.class_embedding = nn.Parameter(scale * torch.randn(width))
        self.positional_embedding = nn.Parameter(scale * torch.randn((resolution_after // patch_size) ** 2 + 1, width))
        self.ln_pre = nn.LayerNorm(width)
        self.transformer = _Transformer(width, layers, heads)
        self.ln_post = nn.LayerNorm(width)


class ClipVisionEncoder(nn.Module):
    def __init__(self, resolution_after, patch_size=16, width=768, layers=12, heads=12, compute_dtype=torch.bfloat16):
        super().__init__()
        if width != 768 or width // heads != 64:
            raise NotImplementedError("the sm_100a kernels are instantiated for ViT-B (width 768, 12 heads x 64)")
        self.visual = _VisualTransformer(patch_size, width, layers, heads, resolution_after)
        self.compute_dtype = compute_dtype
        proj_std = (width ** -0.5) * ((2 * layers) ** -0.5)
        for b in self.visual.transformer.resblocks:  # CLIP.initialize_parameters, clip_model.py:172-180
            nn.init.normal_(b.attn.in_proj_weight, std=width ** -0.5)
            nn.init.normal_(b.attn.out_proj.weight, std=proj_std)
            nn.init.normal_(b.mlp.c_fc.weight, std=(2 * width) ** -0.5)
            nn.init.normal_(b.mlp.c_proj.weight, std=proj_std)

    def forward_patch_embed(self, image):
        """clip_model.py:96-101 (= forward_patch_embed :114-120): conv1 as a GEMM over the 16x16x3 patch matrix, then the
        class token.  -> [B, L, width] in the compute dtype."""
        v = self.visual
        p = v.patch_size
        B, C, Hh, Ww = image.shape
        gh, gw = Hh // p, Ww // p
        # K index = c*p*p + py*p + px: the layout of conv1.weight.view(width, C*p*p)
        patches = image.reshape(B, C, gh, p, gw, p).permute(0, 2, 4, 1, 3, 5).reshape(B * gh * gw, C * p * p)
        w = v.conv1.weight.view(v.conv1.weight.shape[0], -1)
        x = Fn.linear(patches, w, None, dtype=self.compute_dtype).view(B, gh * gw, -1)
        cls = v.class_embedding.to(x.dtype).expand(B, 1, -1)
        return torch.cat([cls, x], dim=1)

    def forward_trans(self, x):
        """clip_model.py:104-110 (= forward_trans :126-132): ln_pre, 11 residual attention blocks, ln_post."""
        v = self.visual
        dt = self.compute_dtype
        B, L, D = x.shape
        x2 = x.reshape(B * L, D)
        if x2.dtype != dt:
            x2 = x2.to(dt)
        x2 = Fn.layernorm(x2, v.ln_pre.weight, v.ln_pre.bias, v.ln_pre.eps)
        for blk in v.transformer.resblocks:
            if VIT_FUSED:
                x2 = vit_block(blk, x2, B, L)   # the same seven kernels as below, as ONE autograd node (vit_block.py)
                continue
            nh = blk.n_head
            h = Fn.layernorm(x2, blk.ln_1.weight, blk.ln_1.bias, blk.ln_1.eps)
            qkv = Fn.linear(h, blk.attn.in_proj_weight, blk.attn.in_proj_bias, dtype=dt)
            a = Fn.attention_packed(qkv, None, B, nh, L, D // nh)
            x2 = Fn.linear(a, blk.attn.out_proj.weight, blk.attn.out_proj.bias, dtype=dt, residual=x2)   # x + attn(ln_1(x))
            h = Fn.layernorm(x2, blk.ln_2.weight, blk.ln_2.bias, blk.ln_2.eps)
            x2 = Fn.ffn(h, blk.mlp.c_fc.weight, blk.mlp.c_fc.bias, blk.mlp.c_proj.weight, blk.mlp.c_proj.bias, residual=x2,
                        act=ops.ACT_QUICKGELU, dtype=dt)                                                  # x + mlp(ln_2(x))
        x2 = Fn.layernorm(x2, v.ln_post.weight, v.ln_post.bias, v.ln_post.eps)
        return x2.view(B, L, D)

    def forward(self, image, image_mask=None):
        if image_mask is not None:
            raise NotImplementedError("key-padding masks on image tokens are not used on the M3AE path (m3ae_module.py:251)")
        if not image.is_cuda:
            raise RuntimeError("ClipVisionEncoder (B200) needs CUDA tensors: there is no CPU fallback")
        x = self.forward_patch_embed(image)
        x = x + self.visual.positional_embedding.to(x.dtype)  # clip_model.py:102
        return self.forward_trans(x)


# ----------------------------------------------------------------------------- RoBERTa-base encoder layers
def _roberta_params(layer):
    """The layer's parameters in functional.PARAM_NAMES order (crossattention.* slots None: a RobertaLayer is a BertCrossLayer
    without the cross-attention block; HF names attention.self.{query,key,value}, attention.output.{dense,LayerNorm},
    intermediate.dense, output.{dense,LayerNorm})."""
    att, so, out = layer.attention.self, layer.attention.output, layer.output
    p = [None] * Fn.NP
    p[Fn.WQ], p[Fn.BQ], p[Fn.WK], p[Fn.BK], p[Fn.WV], p[Fn.BV] = (att.query.weight, att.query.bias, att.key.weight, att.key.bias,
                                                                att.value.weight, att.value.bias)
    p[Fn.WO], p[Fn.BO], p[Fn.G1], p[Fn.B1] = so.dense.weight, so.dense.bias, so.LayerNorm.weight, so.LayerNorm.bias
    p[Fn.W1], p[Fn.BI] = layer.intermediate.dense.weight, layer.intermediate.dense.bias
    p[Fn.W2], p[Fn.BO3], p[Fn.G3], p[Fn.B3] = out.dense.weight, out.dense.bias, out.LayerNorm.weight, out.LayerNorm.bias
    return p


def roberta_layer(layer, h, mask_add, B, L, compute_dtype, training):
    """One HF RobertaLayer on rows h [B*L, 768]: self-attention (RobertaSelfAttention + RobertaSelfOutput) and FFN
    (RobertaIntermediate + RobertaOutput), post-LN, erf GELU — the arithmetic of BertCrossLayer's self-attention and FFN
    blocks (bert_model.py:253-364,416-442) with this layer's own parameters, so it runs as ONE fused autograd node of
    `functional` (LayerConfig.has_cross False: persistent workspace, cached launch descriptors, one flat gradient buffer) instead
    of eight composable nodes per layer: the whole-model step is host-bound and the 12 text layers were a third of its host time
    (tools/whole_step_probe.py).  mask_add: additive key mask [B, L] fp32."""
    att, so = layer.attention.self, layer.attention.output
    key = (compute_dtype, float(att.dropout.p), float(so.dropout.p), float(so.LayerNorm.eps), int(att.num_attention_heads))
    state = layer.__dict__.get("_m3ae_fused")
    if state is None or state[0] != key:
        cfg = Fn.LayerConfig(num_heads=key[4], eps=key[3], hidden_dropout=key[2], attn_dropout=key[1], compute_dtype=compute_dtype,
                             has_cross=False)
        state = (key, cfg, Fn.LayerOperands())
        layer.__dict__["_m3ae_fused"] = state
    _, cfg, operands = state
    D = h.shape[-1]
    stream = dict(hidden=h.view(B, L, D), encoder=None, mask=mask_add, enc_mask=None, params=_roberta_params(layer), operands=operands)
    (out,), = Fn.cross_layers_apply(cfg, training, [stream])
    return out.reshape(B * L, D)


def text_encoder_forward(language_encoder, text_ids, text_masks, compute_dtype, training):
    """m3ae_module.py:230-234: embeddings (HF module: gathers + LayerNorm + dropout), the HF-4.6.0 extended mask, then the
    encoder layers one by one.  -> [B, L, 768] in the compute dtype."""
    emb = language_encoder.embeddings(input_ids=text_ids)
    B, L, D = emb.shape
    mask_add = ((1.0 - text_masks.to(torch.float32)) * -10000.0).contiguous()
    h = emb.reshape(B * L, D).to(compute_dtype)
    for layer in language_encoder.encoder.layer:
        h = roberta_layer(layer, h, mask_add, B, L, compute_dtype, training)
    return h.view(B, L, D)
