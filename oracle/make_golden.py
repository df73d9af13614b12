"""ORACLE — test infrastructure only.  Generates tests/golden/*.pt by running the UNMODIFIED reference modules
(imported from /root/reference through oracle/ref_shim.py) on seeded weights and inputs.

Run in the build container:  python oracle/make_golden.py
The fixtures hold only outputs (inputs/weights are regenerated from seeds by oracle/m3ae_oracle.py helpers), so they
stay small.  tests/test_oracle_golden.py pins oracle/m3ae_oracle.py against them; the GPU parity tests compare the
CUDA path with both.
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import m3ae_oracle as O  # noqa: E402
from oracle import ref_shim  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def probe(shape, seed):
    """Fixed random projection used as the scalar 'loss' for gradient fixtures (a loss that reads only
    post-LayerNorm token-0 energy is degenerate)."""
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed))


def cross_layer_cases():
    # (name, seed, B, Lq, Lk, ragged mask on queries?, ragged mask on keys?)
    return [("text_like", 11, 2, 32, 325, True, False), ("image_like", 12, 1, 325, 32, False, True)]


def cross_layer_inputs(seed, B, Lq, Lk, q_masked, k_masked):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, Lq, 768, generator=g)
    y = torch.randn(B, Lk, 768, generator=g)

    def mk(L, ragged):
        lens = torch.randint(max(1, L // 4), L + 1, (B,), generator=g) if ragged else torch.full((B,), L)
        return (torch.arange(L)[None, :] < lens[:, None]).long()

    return x, y, mk(Lq, q_masked), mk(Lk, k_masked)


def gen_cross_layer(ref):
    out = {}
    for name, seed, B, Lq, Lk, qm, km in cross_layer_cases():
        sd = O.make_cross_layer_state(seed)
        layer = ref.BertCrossLayer(ref_shim.fusion_config()).eval()
        layer.load_state_dict(sd, strict=True)
        x, y, mq, mk = cross_layer_inputs(seed, B, Lq, Lk, qm, km)
        x.requires_grad_(True)
        y.requires_grad_(True)
        r = layer(x, y, O.extended_attention_mask(mq), O.extended_attention_mask(mk), output_attentions=True)
        loss = (r[0] * probe(r[0].shape, seed + 100)).sum()
        loss.backward()
        out[name] = {
            "tok_stride": 1 if Lq <= 32 else 5,  # image-like case keeps every 5th token to stay small
            "layer_out": r[0].detach()[:, :: (1 if Lq <= 32 else 5)].clone(),
            "self_probs_rowsum": r[1].detach().sum(-1), "self_probs_00": r[1].detach()[:, :, 0].clone(),
            "cross_probs_00": r[2].detach()[:, :, 0].clone(),
            "loss": loss.detach().clone(), "dx": x.grad[:, :: (1 if Lq <= 32 else 5)].clone(), "dy_norm": y.grad.norm(),
            "dy_row0": y.grad[:, 0].clone(),
            "param_grad_norms": {k: p.grad.norm() for k, p in layer.named_parameters()},
            "param_grad_samples": {k: p.grad.flatten()[:: max(1, p.numel() // 64)][:64].clone() for k, p in layer.named_parameters()},
        }
    torch.save(out, os.path.join(OUT, "cross_layer.pt"))
    print("cross_layer.pt", {k: float(v["loss"]) for k, v in out.items()})


def gen_fusion():
    mm3, Cls = ref_shim.load_m3ae_module()
    torch.manual_seed(0)
    model = Cls(ref_shim.m3ae_config()).eval()
    sd = O.make_fusion_state(21)
    missing, unexpected = model.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all(k.startswith(("vision_encoder", "language_encoder")) for k in missing), missing
    inp = O.make_fusion_inputs(22, B=2)
    # plug the seeded uni-modal features in; everything from m3ae_module.py:235 on runs as written
    class Const(torch.nn.Module):
        def __init__(self, value):
            super().__init__()
            self.value = value

        def forward(self, *a, **k):
            return self.value

    model.language_encoder.embeddings = Const(inp["text_feats"])
    model.vision_encoder = Const(inp["image_feats"])
    batch = {"image": [torch.zeros(2, 3, 288, 288)], "text_ids": torch.zeros(2, 32, dtype=torch.long),
             "text_labels": torch.full((2, 32), -100), "text_masks": inp["text_masks"]}
    with torch.no_grad():
        ret = model.infer(batch)
        logits = model.vqa_head(ret["multi_modal_cls_feats"])
        loss = torch.nn.functional.binary_cross_entropy_with_logits(logits, inp["targets"]) * inp["targets"].shape[1]
    out = {"multi_modal_cls_feats": ret["multi_modal_cls_feats"], "multi_modal_text_feats": ret["multi_modal_text_feats"],
           "multi_modal_image_feats_tok0": ret["multi_modal_image_feats"][:, 0].clone(),
           "extended_text_masks": ret["extended_text_masks"], "vqa_logits": logits, "vqa_loss": loss,
           "keys": sorted(ret.keys())}
    torch.save(out, os.path.join(OUT, "fusion.pt"))
    print("fusion.pt loss", float(loss), "argmax", logits.argmax(1).tolist())


def gen_full_model():
    """BASELINE config 1 (classifier forward, B=8, 288 px / 32 tokens, CPU fp32, eval) through the reference's REAL
    M3AETransformerSS: its own CLIP ViT-B/16 class + a random-init HF RobertaModel + 6+6 fusion layers + vqa_head, all
    parameters seeded by name (oracle.seed_model_parameters) so the drop-in can rebuild the same weights on the GPU box.
    Also a B=2 forward+backward (eval mode = dropout off) for encoder / fusion gradient parity."""
    mm3, Cls = ref_shim.load_m3ae_module(real_encoders=True)
    torch.manual_seed(0)
    model = Cls(ref_shim.m3ae_config()).eval()
    O.seed_model_parameters(model, 5)
    names = [n for n, _ in model.named_parameters()]
    params = dict(model.named_parameters())
    ref_shim.wrap_text_layers(model)
    cap = {}
    model.vision_encoder.register_forward_hook(lambda m, i, o: cap.__setitem__("uni_image", o.detach()))
    model.multi_modal_language_proj.register_forward_hook(lambda m, i, o: cap.__setitem__("uni_text", i[0].detach()))
    out = {}
    batch = O.make_model_batch(51, 8)
    with torch.no_grad():
        ret = model.infer(batch)
        logits = model.vqa_head(ret["multi_modal_cls_feats"])
        targets = O.vqa_targets(batch["vqa_labels"], batch["vqa_scores"], 498)
        loss = O.vqa_loss(logits, targets)
    out["b8"] = {"vqa_logits": logits, "vqa_loss": loss, "argmax": logits.argmax(1),
                 "multi_modal_cls_feats": ret["multi_modal_cls_feats"], "multi_modal_text_feats": ret["multi_modal_text_feats"],
                 "multi_modal_image_feats_sub": ret["multi_modal_image_feats"][:, ::25].clone(),
                 "uni_text_feats": cap["uni_text"].clone(), "uni_image_feats_sub": cap["uni_image"][:, ::25].clone(),
                 "patched_images_sum": ret["patched_images"].sum(-1), "keys": sorted(ret.keys())}
    batch = O.make_model_batch(52, 2)
    model.zero_grad()
    ret = model.infer(batch)
    logits = model.vqa_head(ret["multi_modal_cls_feats"])
    loss = O.vqa_loss(logits, O.vqa_targets(batch["vqa_labels"], batch["vqa_scores"], 498))
    loss.backward()
    g = {}
    for n in names:
        p = params[n]
        if p.grad is not None:
            g[n] = {"norm": p.grad.norm(), "samples": p.grad.flatten()[:: max(1, p.numel() // 64)][:64].clone()}
    out["b2"] = {"vqa_logits": logits.detach(), "vqa_loss": loss.detach(), "grads": g, "no_grad": [n for n in names if params[n].grad is None]}
    torch.save(out, os.path.join(OUT, "full_model.pt"))
    print("full_model.pt loss b8", float(out["b8"]["vqa_loss"]), "argmax", out["b8"]["argmax"].tolist(), "grads", len(g),
          "without grad", len(out["b2"]["no_grad"]))


DECODER_MODEL_ANSWERS = [["left lower lobe pneumonia"], ["no"]]


def decoder_model_config():
    cfg = ref_shim.m3ae_config()
    cfg.update({"decoder_load_path": "", "mm_encoder_inputs_include_imagetext_feats": True,
                "mm_encoder_inputs_include_cls_feats": True, "learning_rate": 1e-5, "weight_decay": 0.01, "end_lr": 0,
                "decay_power": 1, "optim_type": "adamw", "warmup_steps": 0.1})
    return cfg


def gen_decoder_model():
    """The reference's DecoderModel.forward (m3ae_decoder.py:296-388), eval mode (dropout off), B=2: teacher-forced loss +
    decoder gradients, and the test branch (greedy search_path with the constructor's max_len=128)."""
    tok = O.SynthTokenizer()
    Cls = ref_shim.load_decoder_model(tok)
    torch.manual_seed(0)
    model = Cls(decoder_model_config()).eval()
    O.seed_model_parameters(model, 9)
    ref_shim.wrap_text_layers(model.m3ae)
    batch = O.make_model_batch(61, 2)
    batch["vqa_answer"] = DECODER_MODEL_ANSWERS
    model.current_tasks = ["vqa"]
    ret = model(batch)
    loss = ret["vqa_loss"]
    loss.backward()
    grads = {n: {"norm": p.grad.norm(), "samples": p.grad.flatten()[:: max(1, p.numel() // 64)][:64].clone()}
             for n, p in model.named_parameters() if p.grad is not None}
    with torch.no_grad():
        mem = model.m3ae.infer(batch)
        memory = torch.cat([mem["multi_modal_image_feats"], mem["multi_modal_text_feats"], mem["multi_modal_cls_feats"].view(-1, 2, 768)], dim=1)
        ids = model.decoder.search_path(memory, tok)
        test_ret = model(batch, test=True)
    out = {"loss": loss.detach(), "train_texts": ret["vqa_logits"], "grads": grads, "memory_sub": memory[:, ::20].clone(),
           "greedy_ids": ids, "test_texts": test_ret["vqa_logits"], "test_loss": test_ret["vqa_loss"],
           "ret_keys": sorted(ret.keys())}
    torch.save(out, os.path.join(OUT, "decoder_model.pt"))
    print("decoder_model.pt loss", float(loss), "grads", len(grads), "greedy", ids[0, :8].tolist(), "texts", test_ret["vqa_logits"])


T5_HEAD_ANSWERS = [["left lower lobe pneumonia"], ["no"]]
T5_HEAD_QUESTIONS = ["is there a pleural effusion", "what organ is shown in this image"]


def t5_head_config(imagetext=False):
    """finetune_m3ae.sh:9-16 (cls feature only) or its commented variant :29-36 (image/text features, width 10)."""
    cfg = ref_shim.m3ae_config()
    cfg.update({"load_path_t5": "", "t5_max_length": 12, "mm_encoder_inputs_include_cls_feats": not imagetext,
                "mm_encoder_inputs_include_imagetext_feats": imagetext, "mm_encoder_inputs_mm_feats_width": 10 if imagetext else 0,
                "learning_rate": 1e-5, "weight_decay": 0.01, "end_lr": 0, "decay_power": 1, "optim_type": "adamw", "warmup_steps": 0.1})
    return cfg


def gen_t5_head():
    """The reference's T5VQA_MMEncoderInput (m3ae_t5_mm_encoder_input.py:100-295), eval mode (dropout off), B=2, both input
    variants: prepare_inputs (its per-call random nn.Linear layers draw from the CPU RNG: torch.manual_seed(77) right before),
    T5 encoder, teacher-forced loss, gradients of the unfrozen top-2 encoder / decoder blocks, beam-4 generation."""
    tok = O.SynthT5Tokenizer()
    Cls = ref_shim.load_t5_head(tok)
    out = {}
    for variant in ("cls", "imagetext"):
        torch.manual_seed(0)
        model = Cls(t5_head_config(variant == "imagetext")).eval()
        O.seed_model_parameters(model, 13)
        ref_shim.wrap_text_layers(model.m3ae)
        model.unfreeze_top_layers(2, 2)
        batch = O.make_model_batch(71, 2)
        batch["vqa_answer"] = T5_HEAD_ANSWERS
        batch["text"] = T5_HEAD_QUESTIONS
        model.current_tasks = ["vqa"]
        cfg = model.hparams.m3ae_config
        torch.manual_seed(77)
        with torch.no_grad():
            inp = model.prepare_inputs(batch, cfg["mm_encoder_inputs_include_cls_feats"], cfg["mm_encoder_inputs_include_imagetext_feats"],
                                       cfg["mm_encoder_inputs_mm_feats_width"])
            enc = model.t5.encoder(inputs_embeds=inp["inputs_embeds"], attention_mask=inp["attention_mask"]).last_hidden_state
        torch.manual_seed(77)
        ret = model(batch)
        loss = ret["vqa_loss"]
        loss.backward()
        grads = {n: {"norm": p.grad.norm(), "samples": p.grad.flatten()[:: max(1, p.numel() // 64)][:64].clone()}
                 for n, p in model.named_parameters() if p.grad is not None}
        nz = int((inp["inputs_embeds"].abs().sum(-1) > 0).sum(1).max())
        out[variant] = {"loss": loss.detach(), "inputs_embeds_head": inp["inputs_embeds"][:, :nz + 1].clone(), "nonzero_rows": nz,
                        "attention_mask_sum": inp["attention_mask"].sum(1), "encoder_sub": enc[:, ::32].clone(), "grads": grads,
                        "texts": ret["vqa_logits"], "ret_keys": sorted(ret.keys())}
        print("t5_head", variant, "loss", float(loss), "nonzero rows", nz, "grads", len(grads), "texts", ret["vqa_logits"])
    torch.save(out, os.path.join(OUT, "t5_head.pt"))


class Tok:
    cls_token_id, sep_token_id, eos_token_id, pad_token_id = 101, 102, None, 0


def decoder_inputs(seed, B, T, Lm=359):
    g = torch.Generator().manual_seed(seed)
    mem = torch.randn(B, Lm, 768, generator=g)
    ids = torch.randint(1000, 30000, (B, T), generator=g)
    ids[:, 0] = 101
    lens = torch.randint(2, T + 1, (B,), generator=g)
    pm = torch.arange(T)[None, :] < lens[:, None]
    ids = torch.where(pm, ids, torch.zeros_like(ids))
    return mem, ids, pm


def gen_decoder(ref):
    sd = O.make_decoder_state(31)
    dec = ref.Decoder(6, 768, 8, 3072, 0.1, 16, 30522).eval()
    dec.load_state_dict(sd, strict=True)
    mem, ids, pm = decoder_inputs(32, B=4, T=9)
    with torch.no_grad():
        logits, att = dec(ids, pm, mem)
        greedy = dec.search_path(mem, Tok)
    # greedy with an early [SEP]: bias the vocab projection towards 102 for a second fixture
    sd2 = dict(sd)
    sd2["final_linear.bias"] = sd["final_linear.bias"].clone()
    sd2["final_linear.bias"][102] += 4.0
    dec.load_state_dict(sd2, strict=True)
    with torch.no_grad():
        greedy_sep = dec.search_path(mem, Tok)
    out = {"logits_sub": logits[:, :, ::16].clone(), "argmax": logits.argmax(-1), "greedy": greedy, "greedy_sep": greedy_sep,
           "self_att_l6": att["layer6_dec_self"].clone(), "cross_att_l6_sum": att["layer6_dec_cross"].sum(-1)}
    torch.save(out, os.path.join(OUT, "decoder.pt"))
    print("decoder.pt greedy[0]", greedy[0].tolist(), "greedy_sep[0]", greedy_sep[0].tolist())


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    if len(sys.argv) > 1 and sys.argv[1] == "t5_head":  # one fixture only
        gen_t5_head()
        sys.exit(0)
    ref = ref_shim.load()
    gen_cross_layer(ref)
    gen_decoder(ref)
    gen_fusion()
    gen_full_model()
    gen_decoder_model()
    gen_t5_head()
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)) // 1024, "KiB")
