"""Drop-in `BertCrossLayer` (reference: m3ae/modules/language_encoders/bert_model.py:445-503).

Same constructor argument (an HF Bert/Roberta config), same parameter names and shapes (state_dict-compatible:
`attention.self.query.weight`, ..., `output.LayerNorm.bias`, 26 tensors), same `forward` signature and tuple return.
The sub-modules below are parameter containers with the reference's class names (`BertSelfAttention`,
`BertSelfOutput`, `BertAttention`, `BertIntermediate`, `BertOutput`) so `module.apply(init_weights)`
(m3ae_utils.py:101-109, isinstance checks on nn.Linear / nn.LayerNorm), optimizer grouping by parameter name
(m3ae_utils.py:121-200) and checkpoints keep working; the arithmetic of the whole layer runs as one fused autograd
node over the sm_100a kernels (functional.py).  There is no PyTorch fallback: CPU tensors raise.
"""
import torch
import torch.nn as nn

from . import functional as Fn


class BertSelfAttention(nn.Module):
    """Parameter container for bert_model.py:211-350 (query / key / value projections)."""

    def __init__(self, config):
        super().__init__()
        if config.hidden_size % config.num_attention_heads != 0 and not hasattr(config, "embedding_size"):
            raise ValueError(
                f"The hidden size ({config.hidden_size}) is not a multiple of the number of attention "
                f"heads ({config.num_attention_heads})")
        self.num_attention_heads = config.num_attention_heads
        self.attention_head_size = config.hidden_size // config.num_attention_heads
        self.all_head_size = self.num_attention_heads * self.attention_head_size
        self.query = nn.Linear(config.hidden_size, self.all_head_size)
        self.key = nn.Linear(config.hidden_size, self.all_head_size)
        self.value = nn.Linear(config.hidden_size, self.all_head_size)
        self.dropout = nn.Dropout(config.attention_probs_dropout_prob)
        if getattr(config, "position_embedding_type", "absolute") not in (None, "absolute"):
            raise NotImplementedError("only absolute position embeddings are on the M3AE fusion path")


class BertSelfOutput(nn.Module):
    """Parameter container for bert_model.py:353-364 / :431-442 (dense + LayerNorm)."""

    def __init__(self, in_features, config):
        super().__init__()
        self.dense = nn.Linear(in_features, config.hidden_size)
        self.LayerNorm = nn.LayerNorm(config.hidden_size, eps=config.layer_norm_eps)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)


class BertAttention(nn.Module):
    """bert_model.py:367-413."""

    def __init__(self, config):
        super().__init__()
        self.self = BertSelfAttention(config)
        self.output = BertSelfOutput(config.hidden_size, config)
        self.pruned_heads = set()


class BertIntermediate(nn.Module):
    """bert_model.py:416-428 (dense + exact erf GELU)."""

    def __init__(self, config):
        super().__init__()
        self.dense = nn.Linear(config.hidden_size, config.intermediate_size)
        act = config.hidden_act
        if not (act == "gelu" or getattr(act, "__name__", "") == "gelu"):
            raise NotImplementedError(f"hidden_act={act!r}: the fused FFN epilogue implements the exact erf GELU the M3AE config uses")


class BertOutput(BertSelfOutput):
    def __init__(self, config):
        super().__init__(config.intermediate_size, config)


class BertCrossLayer(nn.Module):
    """forward(hidden_states[B,Lq,H], encoder_hidden_states[B,Lk,H], attention_mask[B,1,1,Lq],
    encoder_attention_mask[B,1,1,Lk], output_attentions=False)
      -> (layer_output,) + ((self_probs[B,nh,Lq,Lq], cross_probs[B,nh,Lq,Lk]) if output_attentions)

    compute_dtype: torch.bfloat16 (tcgen05 tensor-core path, default) or torch.float32 (strict-parity path).
    Inputs may be fp32 or bf16; the output has the input's dtype.
    """

    def __init__(self, config, compute_dtype=torch.bfloat16):
        super().__init__()
        if config.hidden_size != 768 or config.intermediate_size % 256 != 0:
            raise NotImplementedError("the sm_100a kernels are instantiated for hidden_size 768 (ViT-B/16 + RoBERTa-base)")
        self.chunk_size_feed_forward = config.chunk_size_feed_forward  # chunking changes nothing numerically (:493-495)
        self.seq_len_dim = 1
        self.attention = BertAttention(config)
        self.is_decoder = config.is_decoder
        self.add_cross_attention = config.add_cross_attention
        self.crossattention = BertAttention(config)
        self.intermediate = BertIntermediate(config)
        self.output = BertOutput(config)
        self.layer_cfg = Fn.LayerConfig(
            num_heads=config.num_attention_heads, eps=config.layer_norm_eps,
            hidden_dropout=config.hidden_dropout_prob, attn_dropout=config.attention_probs_dropout_prob,
            compute_dtype=compute_dtype)
        self._operands = Fn.LayerOperands()

    def set_compute_dtype(self, dtype):
        self.layer_cfg.compute_dtype = dtype
        return self

    def ordered_params(self):
        """The 26 parameters in functional.PARAM_NAMES order."""
        a, c = self.attention, self.crossattention
        return [
            a.self.query.weight, a.self.query.bias, a.self.key.weight, a.self.key.bias, a.self.value.weight, a.self.value.bias,
            a.output.dense.weight, a.output.dense.bias, a.output.LayerNorm.weight, a.output.LayerNorm.bias,
            c.self.query.weight, c.self.query.bias, c.self.key.weight, c.self.key.bias, c.self.value.weight, c.self.value.bias,
            c.output.dense.weight, c.output.dense.bias, c.output.LayerNorm.weight, c.output.LayerNorm.bias,
            self.intermediate.dense.weight, self.intermediate.dense.bias,
            self.output.dense.weight, self.output.dense.bias, self.output.LayerNorm.weight, self.output.LayerNorm.bias,
        ]

    def stream(self, hidden_states, encoder_hidden_states, attention_mask, encoder_attention_mask):
        return dict(hidden=hidden_states, encoder=encoder_hidden_states, mask=attention_mask,
                    enc_mask=encoder_attention_mask, params=self.ordered_params(), operands=self._operands)

    def forward(self, hidden_states, encoder_hidden_states, attention_mask=None, encoder_attention_mask=None,
                output_attentions=False):
        (out,) = Fn.cross_layers_apply(self.layer_cfg, self.training,
                                       [self.stream(hidden_states, encoder_hidden_states, attention_mask, encoder_attention_mask)],
                                       want_probs=output_attentions)
        return out


def fused_layer_pair(text_layer, image_layer, x, y, ext_text_mask, ext_image_mask, output_attentions=False):
    """One fusion index: x1 = text_layer(x, y, ...), y1 = image_layer(y, x, ...) (m3ae_module.py:276-277) issued as a
    single co-scheduled kernel sequence.  Returns (x1_tuple, y1_tuple) exactly as the two separate calls would."""
    cfg = text_layer.layer_cfg
    outs = Fn.cross_layers_apply(cfg, text_layer.training,
                                 [text_layer.stream(x, y, ext_text_mask, ext_image_mask),
                                  image_layer.stream(y, x, ext_image_mask, ext_text_mask)],
                                 want_probs=output_attentions)
    return outs[0], outs[1]
