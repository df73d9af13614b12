"""ORACLE — test infrastructure only.  Import shim that loads the UNMODIFIED reference modules from /root/reference.

Only usable where /root/reference exists (the build container, not the GPU box): used by oracle/make_golden.py to
produce tests/golden/*.pt and by tests that pin oracle/m3ae_oracle.py against the real reference.

Why a shim is needed:
  - m3ae/modules/__init__.py:4 imports a file that does not exist -> register the package without running __init__;
  - bert_model.py:48-53 imports helpers that moved out of transformers.modeling_utils after the pinned 4.6.0;
  - pytorch_lightning / torchmetrics / timm are absent -> stub modules.
Nothing from the reference is copied: its files are imported in place.
"""
import importlib
import os
import sys
import types

REF_ROOT = os.environ.get("M3AE_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REF_ROOT, "m3ae", "modules"))


def _install_stubs():
    import torch.nn as nn
    import transformers
    import transformers.modeling_utils as mu
    import transformers.pytorch_utils as pu

    for name in ("apply_chunking_to_forward", "prune_linear_layer"):
        if not hasattr(mu, name):
            setattr(mu, name, getattr(pu, name))
    if not hasattr(mu, "find_pruneable_heads_and_indices"):
        mu.find_pruneable_heads_and_indices = lambda *a, **k: (_ for _ in ()).throw(NotImplementedError())
    fu = sys.modules.get("transformers.file_utils") or importlib.import_module("transformers.file_utils")

    def _noop_decorator(*a, **k):
        def deco(fn):
            return fn
        return deco

    for name in ("add_code_sample_docstrings", "add_start_docstrings", "add_start_docstrings_to_model_forward",
                 "replace_return_docstrings"):
        setattr(fu, name, _noop_decorator)
    if not hasattr(fu, "ModelOutput"):
        from transformers.utils import ModelOutput
        fu.ModelOutput = ModelOutput
    import transformers.optimization as opt
    if not hasattr(opt, "AdamW"):
        import torch
        opt.AdamW = torch.optim.AdamW
    if not hasattr(transformers, "AdamW"):
        import torch
        transformers.AdamW = torch.optim.AdamW

    if "pytorch_lightning" not in sys.modules:
        pl = types.ModuleType("pytorch_lightning")

        class LightningModule(nn.Module):
            def save_hyperparameters(self, *a, **k):
                import inspect
                frame = inspect.currentframe().f_back
                args = {k_: v for k_, v in frame.f_locals.items() if k_ not in ("self", "__class__")}
                self.hparams = types.SimpleNamespace(**args)

            def log(self, *a, **k):
                pass

            @property
            def device(self):
                return next(self.parameters()).device

        pl.LightningModule = LightningModule
        sys.modules["pytorch_lightning"] = pl


def _pkg(name, path):
    m = types.ModuleType(name)
    m.__path__ = [path]
    sys.modules[name] = m
    return m


def load():
    """Returns a namespace with the reference classes on the hot path: BertCrossLayer, Decoder, DecoderLayer, Pooler."""
    if not available():
        raise RuntimeError(f"reference tree not found at {REF_ROOT}; the shim only works in the build container")
    _install_stubs()
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    if "m3ae" not in sys.modules or not hasattr(sys.modules["m3ae"], "__path__"):
        _pkg("m3ae", os.path.join(REF_ROOT, "m3ae"))
    # bypass the broken m3ae/modules/__init__.py (imports a missing file)
    mods = _pkg("m3ae.modules", os.path.join(REF_ROOT, "m3ae", "modules"))
    _pkg("m3ae.modules.language_encoders", os.path.join(REF_ROOT, "m3ae", "modules", "language_encoders"))
    bert = importlib.import_module("m3ae.modules.language_encoders.bert_model")
    # m3ae_decoder.py imports M3AETransformerSS / m3ae_t5_utils / objectives from the package: give it placeholders,
    # the Decoder / DecoderLayer classes need only torch.
    mods.M3AETransformerSS = type("M3AETransformerSS", (), {})
    for sub in ("m3ae_t5_utils", "objectives"):
        if f"m3ae.modules.{sub}" not in sys.modules:
            stub = types.ModuleType(f"m3ae.modules.{sub}")
            sys.modules[f"m3ae.modules.{sub}"] = stub
            setattr(mods, sub, stub)
    dec = importlib.import_module("m3ae.modules.m3ae_decoder")
    ns = types.SimpleNamespace(BertCrossLayer=bert.BertCrossLayer, BertSelfAttention=bert.BertSelfAttention,
                               Decoder=dec.Decoder, DecoderLayer=dec.DecoderLayer, bert_model=bert, m3ae_decoder=dec)
    return ns


def fusion_config(drop_rate=0.0):
    """The RobertaConfig M3AETransformerSS builds for its fusion layers (m3ae_module.py:24-33) with the
    task_finetune_vqa_vqa_rad + clip16 + text_roberta values (config.py:45-53,176-198,245-259)."""
    from transformers import RobertaConfig
    return RobertaConfig(vocab_size=50265, hidden_size=768, num_hidden_layers=6, num_attention_heads=12,
                         intermediate_size=768 * 4, max_position_embeddings=32, hidden_dropout_prob=drop_rate,
                         attention_probs_dropout_prob=drop_rate)
