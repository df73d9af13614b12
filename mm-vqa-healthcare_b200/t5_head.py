"""Drop-in for the reference's `T5VQA_MMEncoderInput` (m3ae/modules/m3ae_t5_mm_encoder_input.py:12-460; SURVEY §8 row T1,
BASELINE config 4): the T5 generative answer head — frozen M3AE -> projected multimodal features as T5 ENCODER INPUT
EMBEDDINGS -> teacher-forced T5 loss (train / val) or beam-4 generation (test).

What runs where.  The frozen `m3ae.infer(batch)` (103 GF/sample, the whole cost of this head besides T5-small's 26 GF) runs
on this repo's sm_100a kernels (`M3AETransformerSS`).  The T5 arithmetic is third-party code in the reference too (HF
`T5ForConditionalGeneration`, transformers==4.6.0 pinned at requirements.txt:4) and stays the installed HF implementation
(library code: cuBLAS / SDPA under bf16 autocast) — SURVEY §8(f)#3 (native T5 kernels) is not built.  Parity of this head is
"unpinned" in the reference (no test, no checkpoint); here it is pinned to outputs of the reference's own class run through
oracle/ref_shim.py with a random-init t5-small (tests/golden/t5_head.pt, tests/test_t5_head_gpu.py).

forward(batch, test=False) follows m3ae_t5_mm_encoder_input.py:100-295:
  1. `prepare_inputs` (:100-188): per sample, the embeddings of the prompt "question:" (T5's shared embedding), then — as
     configured — the [1536] cls feature through a FRESH random `nn.Linear(1536, 512)` and the image / text token matrices
     through fresh random `nn.Linear(L, width/2)` (over the sequence axis) and `nn.Linear(768, 512)`; zero-padded to 512
     positions; attention mask all ones (so the padding IS attended, :174-178).  The reference constructs those Linear layers
     anew for every sample of every call (:75-77, :128-145) on the CPU and moves them to the device: they draw from the CPU RNG.
     This class constructs the same layers in the same order (same RNG consumption => same numbers under the same seed) but
     applies them as ONE batched matmul per projection instead of 5 tiny launches per sample.
  2. `t5.encoder(inputs_embeds, attention_mask)` (:198-201).
  3. test / no task: `t5.generate(num_beams=4, ...)` (:205-214).  train / val: tokenise the first answer of every sample,
     `t5(encoder_outputs=..., labels=...)` -> loss (:231-244), then — for the metrics — the same beam-4 generation under
     no_grad (:250-259; `generate_for_metrics=False` skips it: BASELINE config 4 is "teacher-forced").
  4. result dict of objectives.compute_vqa (objectives.py:122-147).

`T5Tokenizer.from_pretrained('t5-small')` / `T5ForConditionalGeneration.from_pretrained('t5-small')` (:26-27) need files that
do not exist offline: pass `tokenizer=` / `t5=`, or the class falls back to `HashT5Tokenizer` and a random-init t5-small.
"""
import torch
import torch.nn as nn

from .m3ae_module import M3AETransformerSS, _Base

MAX_SEQ_LEN = 512  # m3ae_t5_mm_encoder_input.py:160


class HashT5Tokenizer:
    """Whitespace tokenizer with T5's special ids (pad 0, eos 1) and hashed ids in [3, 32000): stands in for T5Tokenizer when the
    sentencepiece model is unavailable (offline boxes, tests, benchmarks).  Same surface the head uses."""

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


def _default_tokenizer():
    try:
        from transformers import T5Tokenizer
        return T5Tokenizer.from_pretrained("t5-small")  # m3ae_t5_mm_encoder_input.py:26
    except Exception:
        return HashT5Tokenizer()


def _default_t5():
    from transformers import T5Config, T5ForConditionalGeneration
    try:
        return T5ForConditionalGeneration.from_pretrained("t5-small")  # :27
    except Exception:
        return T5ForConditionalGeneration(T5Config(decoder_start_token_id=0))  # t5-small's architecture, random weights


class T5VQA_MMEncoderInput(_Base):
    def __init__(self, m3ae_config, freeze_m3ae=True, freeze_t5_layers=True, tokenizer=None, t5=None, compute_dtype=None,
                 generate_for_metrics=True):
        super().__init__()
        self.hparams = type("HParams", (), {})()
        self.hparams.m3ae_config = m3ae_config
        self.m3ae = M3AETransformerSS(m3ae_config, compute_dtype)
        if freeze_m3ae:  # :22-24
            for p in self.m3ae.parameters():
                p.requires_grad = False
        self.tokenizer = tokenizer if tokenizer is not None else _default_tokenizer()
        self.t5 = t5 if t5 is not None else _default_t5()
        path = m3ae_config.get("load_path_t5", "")
        if path:  # :29-32
            self.load_state_dict(torch.load(path, map_location="cpu")["state_dict"], strict=False)
        if freeze_t5_layers:  # :35-37
            for p in self.t5.parameters():
                p.requires_grad = False
        # constructed by the reference (:40-43) and never used in its forward; kept for state-dict compatibility
        self.feature_projection = nn.Linear(m3ae_config["hidden_size"] * 2, self.t5.config.hidden_size)
        self.max_answer_length = m3ae_config["t5_max_length"]
        self.generate_for_metrics = generate_for_metrics
        self.current_tasks = list()

    # ---- m3ae_t5_mm_encoder_input.py:75-98
    def projection_layer(self, input_dim, output_dim=512):
        return nn.Linear(input_dim, output_dim)  # CPU construction = the reference's RNG consumption; moved in batches below

    def unfreeze_top_layers(self, num_encoder_layers=2, num_decoder_layers=2):
        for p in self.t5.parameters():
            p.requires_grad = False
        enc, dec = self.t5.encoder.block, self.t5.decoder.block
        for i in range(len(enc) - num_encoder_layers, len(enc)):
            for p in enc[i].parameters():
                p.requires_grad = True
        for i in range(len(dec) - num_decoder_layers, len(dec)):
            for j in (0, 1):  # self attention, cross attention (:91-96; the decoder FFN stays frozen, as in the reference)
                for p in dec[i].layer[j].parameters():
                    p.requires_grad = True

    # ---- :100-188
    def prepare_inputs(self, batch, include_cls_feats, include_imagetext_feats, mm_feats_width):
        out = self.m3ae.infer(batch, mask_text=False, mask_image=False)
        dev = out["multi_modal_cls_feats"].device
        B = len(batch["text"])
        half = mm_feats_width // 2
        cls = out["multi_modal_cls_feats"].float() if include_cls_feats else None
        img = out["multi_modal_image_feats"].float() if include_imagetext_feats else None
        txt = out["multi_modal_text_feats"].float() if include_imagetext_feats else None
        # the reference's per-sample layer constructions, in its order (cls; image seq, image embed; text seq, text embed)
        layers = {k: [] for k in ("cls", "img_seq", "img_emb", "txt_seq", "txt_emb")}
        for _ in range(B):
            if include_cls_feats:
                layers["cls"].append(self.projection_layer(cls.shape[1]))
            if include_imagetext_feats:
                layers["img_seq"].append(self.projection_layer(img.shape[1], half))
                layers["img_emb"].append(self.projection_layer(img.shape[2]))
                layers["txt_seq"].append(self.projection_layer(txt.shape[1], half))
                layers["txt_emb"].append(self.projection_layer(txt.shape[2]))

        def stacked(name):
            w = torch.stack([l.weight.detach() for l in layers[name]]).to(dev, non_blocking=True)
            b = torch.stack([l.bias.detach() for l in layers[name]]).to(dev, non_blocking=True)
            return w, b

        prefix = self.tokenizer("question:", return_tensors="pt", add_special_tokens=False).input_ids.to(dev)  # :112-117
        parts = [self.t5.shared(prefix).expand(B, -1, -1).float()]
        if include_cls_feats:  # :121-125
            w, b = stacked("cls")
            parts.append((torch.bmm(cls.unsqueeze(1), w.transpose(1, 2)) + b.unsqueeze(1)))
        if include_imagetext_feats:  # :130-146: Linear over the sequence axis, then over the embedding axis
            for feats, seq, emb in ((img, "img_seq", "img_emb"), (txt, "txt_seq", "txt_emb")):
                ws, bs = stacked(seq)
                we, be = stacked(emb)
                red = torch.bmm(ws, feats) + bs.unsqueeze(2)                      # [B, half, 768] = (W_seq · feats) + b_seq per row
                parts.append(torch.bmm(red, we.transpose(1, 2)) + be.unsqueeze(1))  # [B, half, 512]
        emb = torch.cat(parts, dim=1)
        L = emb.shape[1]
        if L < MAX_SEQ_LEN:  # :162-171
            emb = torch.nn.functional.pad(emb, (0, 0, 0, MAX_SEQ_LEN - L))
        elif L > MAX_SEQ_LEN:
            emb = emb[:, :MAX_SEQ_LEN]
        mask = torch.ones(B, emb.shape[1], dtype=torch.long, device=dev)  # :174-178: padding is attended
        return {"inputs_embeds": emb, "attention_mask": mask}

    def _autocast(self, dev):
        on = self.m3ae.compute_dtype == torch.bfloat16 and dev.type == "cuda"
        return torch.autocast("cuda", dtype=torch.bfloat16, enabled=on)

    def _generate(self, encoder_outputs, **extra):
        return self.t5.generate(encoder_outputs=encoder_outputs, max_length=self.max_answer_length, num_beams=4, early_stopping=True,
                                pad_token_id=self.tokenizer.pad_token_id, eos_token_id=self.tokenizer.eos_token_id,
                                return_dict_in_generate=True, **extra)

    # ---- :191-295
    def forward(self, batch, test=False):
        cfg = self.hparams.m3ae_config
        inp = self.prepare_inputs(batch, cfg["mm_encoder_inputs_include_cls_feats"], cfg["mm_encoder_inputs_include_imagetext_feats"],
                                  cfg["mm_encoder_inputs_mm_feats_width"])
        dev = inp["inputs_embeds"].device
        ret = {}
        with self._autocast(dev):
            encoder_outputs = self.t5.encoder(inputs_embeds=inp["inputs_embeds"], attention_mask=inp["attention_mask"])
            if len(self.current_tasks) == 0 or test:
                with torch.no_grad():
                    outputs = self._generate(encoder_outputs, output_scores=True)
                ret["generated_texts"] = [[t] for t in self.tokenizer.batch_decode(outputs.sequences, skip_special_tokens=True)]
                return ret  # the reference returns the empty dict here (:222-223); the texts are an addition
            labels = batch["vqa_answer"]
            flat = [a[0] for a in labels]
            label_tokens = self.tokenizer(flat, padding=True, truncation=True, return_tensors="pt").input_ids.to(dev)
            outputs = self.t5(encoder_outputs=encoder_outputs, labels=label_tokens, return_dict=True)
            texts = [[""] for _ in flat]
            if self.generate_for_metrics:
                with torch.no_grad():
                    gen = self._generate(encoder_outputs)
                texts = [[t] for t in self.tokenizer.batch_decode(gen.sequences, skip_special_tokens=True)]
        loss = outputs.loss.float()
        phase = "test" if test else ("train" if self.training else "val")
        self.log(f"vqa/{phase}/loss", loss.detach())
        ret.update({"vqa_loss": loss, "vqa_logits": texts, "vqa_labels": labels, "vqa_scores": batch.get("vqa_scores"),
                    "vqa_answer_types": torch.tensor(batch.get("answer_types", [0] * len(flat)), device=dev),
                    "t5_logits": outputs.logits})
        return ret

    # ---- pytorch_lightning hooks (:336-372)
    def _set_task(self):
        self.current_tasks = [k for k, v in self.hparams.m3ae_config["loss_names"].items() if v > 0]

    def training_step(self, batch, batch_idx):
        self._set_task()
        output = self(batch)
        total = sum(v * self.hparams.m3ae_config["loss_names"][k.replace("_loss", "")] for k, v in output.items() if k.endswith("_loss"))
        return {"loss": total}

    def validation_step(self, batch, batch_idx):
        self._set_task()
        return self(batch)

    def test_step(self, batch, batch_idx):
        self._set_task()
        return self(batch, test=True)

    def training_epoch_end(self, outs):
        pass

    validation_epoch_end = test_epoch_end = training_epoch_end

    def configure_optimizers(self, max_steps=1000):
        """m3ae_t5_utils.set_schedule (:130-288) reduced to what this head trains: AdamW(eps 1e-8, betas (0.9, 0.98)) over the
        parameters that require grad, weight decay except bias / norm weights, polynomial decay with warm-up."""
        c = self.hparams.m3ae_config
        lr, wd = c.get("learning_rate", 1e-5), c.get("weight_decay", 0.01)
        no_decay = ["bias", "LayerNorm.bias", "LayerNorm.weight", "norm.bias", "norm.weight", "layer_norm.weight"]
        named = [(n, p) for n, p in self.named_parameters() if p.requires_grad]
        groups = [{"params": [p for n, p in named if not any(nd in n for nd in no_decay)], "weight_decay": wd, "lr": lr},
                  {"params": [p for n, p in named if any(nd in n for nd in no_decay)], "weight_decay": 0.0, "lr": lr}]
        opt = torch.optim.AdamW([g for g in groups if g["params"]], lr=lr, eps=1e-8, betas=(0.9, 0.98))
        max_steps = c.get("max_steps") or max_steps
        warm = c.get("warmup_steps", 0.1)
        warm = int(max_steps * warm) if isinstance(warm, float) else warm
        end_lr, power = c.get("end_lr", 0), c.get("decay_power", 1)

        def lr_lambda(step):
            if step < warm:
                return step / max(1, warm)
            if step > max_steps:
                return end_lr / lr
            rem = 1 - (step - warm) / max(1, max_steps - warm)
            return ((lr - end_lr) * rem ** power + end_lr) / lr

        return [opt], [{"scheduler": torch.optim.lr_scheduler.LambdaLR(opt, lr_lambda), "interval": "step"}]
