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

_SNAPSHOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")  # oracle/snapshot_ref.py (git-ignored)
REF_ROOT = os.environ.get("M3AE_REFERENCE_ROOT") or ("/root/reference" if os.path.isdir("/root/reference/m3ae/modules") else _SNAPSHOT)


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


def load_m3ae_module(real_encoders=False):
    """Imports the reference's m3ae/modules/m3ae_module.py unmodified and returns (module, M3AETransformerSS).
    real_encoders=True: the uni-modal encoders are the reference's own CLIP class (vision_encoders/clip_model.py, ViT-B/16
    hyper-parameters that build_model would read from the checkpoint, :277-297) and a random-init HF RobertaModel of
    roberta-base dimensions, with the two HF 4.6.0 -> 5.x drifts neutralised: get_extended_attention_mask
    keeps the pinned (1 - mask) * -10000 semantics, and encoder layers return 1-tuples so that `layer(h, mask)[0]`
    (m3ae_module.py:234) still selects the hidden states.
    The uni-modal encoders need downloaded checkpoints (clip_model.build_model, RobertaModel.from_pretrained) that do
    not exist offline, so both constructors are replaced by tiny stand-ins; callers then plug seeded uni-modal
    features in (see oracle/make_golden.py) so that the fusion section of infer() (m3ae_module.py:235,252-297) runs
    exactly as written."""
    ns = load()
    import torch
    import torch.nn as nn
    mods = sys.modules["m3ae.modules"]
    base = os.path.join(REF_ROOT, "m3ae", "modules")
    # stubs for pieces that need absent third-party packages (torchmetrics, timm)
    if "m3ae.gadgets" not in sys.modules:
        _pkg("m3ae.gadgets", os.path.join(REF_ROOT, "m3ae", "gadgets"))
        mm = types.ModuleType("m3ae.gadgets.my_metrics")
        for n in ("Accuracy", "Scalar", "VQAScore", "VQARADScore", "VQAExactMatch", "ROUGE1Score", "ROUGE2Score",
                  "BLEUScore", "ROCScore", "F1Score"):
            setattr(mm, n, type(n, (nn.Module,), {"forward": lambda self, *a, **k: torch.zeros(())}))
        sys.modules["m3ae.gadgets.my_metrics"] = mm
    _pkg("m3ae.modules.vision_encoders", os.path.join(base, "vision_encoders"))
    for sub in ("swin_transformer", "swin_helpers"):
        stub = types.ModuleType(f"m3ae.modules.vision_encoders.{sub}")
        stub.swin_adapt_position_encoding = lambda *a, **k: None
        sys.modules[f"m3ae.modules.vision_encoders.{sub}"] = stub
        setattr(sys.modules["m3ae.modules.vision_encoders"], sub, stub)
    # real objectives / utils / heads (drop the placeholders load() registered for the decoder import)
    for sub in ("objectives", "m3ae_t5_utils"):
        sys.modules.pop(f"m3ae.modules.{sub}", None)
    for sub in ("m3ae_utils", "objectives", "prediction_heads"):
        m = importlib.import_module(f"m3ae.modules.{sub}")
        setattr(mods, sub, m)
    mm3 = importlib.import_module("m3ae.modules.m3ae_module")

    class _TinyVision(nn.Module):
        def __init__(self):
            super().__init__()
            self.dummy = nn.Parameter(torch.zeros(1))

    class _TinyText(nn.Module):
        def __init__(self):
            super().__init__()
            self.embeddings = nn.Identity()
            self.encoder = nn.Module()
            self.encoder.layer = nn.ModuleList()

        def get_extended_attention_mask(self, mask, shape, device):  # HF 4.6.0 semantics (pinned version)
            return (1.0 - mask[:, None, None, :].to(torch.float32)) * -10000.0

    if not real_encoders:
        mm3.build_model = lambda *a, **k: _TinyVision()
        mm3.RobertaModel = types.SimpleNamespace(from_pretrained=lambda *a, **k: _TinyText())
        return mm3, mm3.M3AETransformerSS

    clip_model = importlib.import_module("m3ae.modules.vision_encoders.clip_model")

    def build_clip(name, resolution_after=224, jit=False):
        return clip_model.CLIP(512, 224, 12, 768, 16, 77, 49408, 512, 8, 12, resolution_after)

    class _TupleLayer(nn.Module):
        """HF 5.x encoder layers return a Tensor; the reference indexes [0] (HF 4.6.0 returned a tuple)."""

        def __init__(self, layer):
            super().__init__()
            self.layer = layer

        def forward(self, h, mask):
            out = self.layer(h, mask)
            return out if isinstance(out, (tuple, list)) else (out,)

    def build_roberta(name, *a, **k):
        from transformers import RobertaConfig, RobertaModel
        rc = RobertaConfig(vocab_size=50265, hidden_size=768, num_hidden_layers=12, num_attention_heads=12,
                           intermediate_size=3072, max_position_embeddings=514, type_vocab_size=1, pad_token_id=1,
                           layer_norm_eps=1e-5)
        try:
            rc._attn_implementation = "eager"
        except Exception:
            pass
        m = RobertaModel(rc)
        m.get_extended_attention_mask = lambda mask, shape, device=None: (1.0 - mask[:, None, None, :].to(torch.float32)) * -10000.0
        return m

    mm3.build_model = build_clip
    mm3.RobertaModel = types.SimpleNamespace(from_pretrained=build_roberta)
    mm3._TupleLayer = _TupleLayer
    return mm3, mm3.M3AETransformerSS


def wrap_text_layers(model):
    """Call after loading weights: wraps language_encoder.encoder.layer[i] so that `layer(h, mask)[0]` works on HF 5.x.
    (The wrappers are not registered under new names in state_dict: they are applied on a shallow ModuleList copy.)"""
    import torch.nn as nn
    mm3 = sys.modules["m3ae.modules.m3ae_module"]
    enc = model.language_encoder.encoder
    object.__setattr__(enc, "_orig_layers", enc.layer)
    enc.layer = nn.ModuleList([mm3._TupleLayer(l) for l in enc._orig_layers])
    return model


def m3ae_config(drop_rate=0.0, image_size=288):
    """config.py defaults (:19-119) + task_finetune_vqa_vqa_rad (:176-198) + clip16 (:245-251) + text_roberta (:256-259)."""
    return {
        "vit": "ViT-B/16", "tokenizer": "roberta-base", "vocab_size": 50265, "hidden_size": 768, "num_layers": 6,
        "num_heads": 12, "mlp_ratio": 4, "max_text_len": 32, "drop_rate": drop_rate, "image_size": image_size,
        "patch_size": 16, "input_text_embed_size": 768, "input_image_embed_size": 768, "num_top_layer": 6,
        "loss_names": {"mlm": 0, "mim": 0, "itm": 0, "vqa": 1, "cls": 0, "irtr": 0}, "load_path": "", "test_only": False,
        "vqa_label_size": 498, "mim_prob": 0.75, "mim_layer": 3,
    }


def load_decoder_model(tokenizer):
    """Imports the reference's DecoderModel (m3ae_decoder.py:185-443) unmodified, wired to the real M3AETransformerSS class
    (random-init real encoders, see load_m3ae_module) and to `tokenizer` in place of BertTokenizer.from_pretrained (:191;
    the vocabulary file does not exist offline).  Returns the class."""
    mm3, Cls = load_m3ae_module(real_encoders=True)
    mods = sys.modules["m3ae.modules"]
    mods.M3AETransformerSS = Cls
    for sub in ("m3ae_t5_utils",):
        sys.modules.pop(f"m3ae.modules.{sub}", None)
        m = importlib.import_module(f"m3ae.modules.{sub}")
        setattr(mods, sub, m)
    sys.modules.pop("m3ae.modules.m3ae_decoder", None)
    dec = importlib.import_module("m3ae.modules.m3ae_decoder")
    dec.BertTokenizer = types.SimpleNamespace(from_pretrained=lambda *a, **k: tokenizer)
    return dec.DecoderModel


def random_t5():
    """t5-small's architecture (T5Config() defaults = t5-small) with random weights: the checkpoint does not exist offline."""
    from transformers import T5Config, T5ForConditionalGeneration
    return T5ForConditionalGeneration(T5Config(decoder_start_token_id=0))


def load_t5_head(tokenizer):
    """Imports the reference's T5VQA_MMEncoderInput (m3ae_t5_mm_encoder_input.py:12-295) unmodified, wired to the real
    M3AETransformerSS class (random-init real encoders), to `tokenizer` in place of T5Tokenizer.from_pretrained('t5-small') (:26)
    and to a random-init t5-small in place of T5ForConditionalGeneration.from_pretrained('t5-small') (:27).  Returns the class."""
    mm3, Cls = load_m3ae_module(real_encoders=True)
    mods = sys.modules["m3ae.modules"]
    mods.M3AETransformerSS = Cls
    for sub in ("m3ae_t5_utils", "objectives"):
        m = sys.modules.get(f"m3ae.modules.{sub}") or importlib.import_module(f"m3ae.modules.{sub}")
        setattr(mods, sub, m)
    sys.modules.pop("m3ae.modules.m3ae_t5_mm_encoder_input", None)
    t5m = importlib.import_module("m3ae.modules.m3ae_t5_mm_encoder_input")
    t5m.T5Tokenizer = types.SimpleNamespace(from_pretrained=lambda *a, **k: tokenizer)
    t5m.T5ForConditionalGeneration = types.SimpleNamespace(from_pretrained=lambda *a, **k: random_t5())
    return t5m.T5VQA_MMEncoderInput
