"""Drop-in for the reference's `DecoderModel` (m3ae/modules/m3ae_decoder.py:185-443): the generative answer head's
LightningModule — frozen M3AE -> multimodal memory -> teacher-forced decoder loss (train / val) or greedy decoding (test).

forward(batch, test=False) follows m3ae_decoder.py:296-388 line by line:
  1. `m3ae.infer(batch)` (:299).  The M3AE is frozen (requires_grad False, :214-216) but stays in whatever mode the
     module is in — in `.train()` its dropouts are active during decoder fine-tuning, exactly as in the reference.
  2. memory = cat([multi_modal_image_feats, multi_modal_text_feats] if `mm_encoder_inputs_include_imagetext_feats`,
     + [multi_modal_cls_feats.view(-1, 2, 768)] if `mm_encoder_inputs_include_cls_feats`, dim=1)  (:302-313)
     -> [B, Li + Lt + 2, 768] = [B, 359, 768] at 288 px.
  3. test (or no current task): `decoder.search_path(memory, tokenizer)` -> `tokenizer.decode(..., skip_special_tokens)`
     (:321-330), loss 0.
  4. train / val: tokenise the first answer of every sample (max_length = max_answer_length, :336-342);
     `target_shifted = tokens[:, :-1]` with [SEP] -> [PAD] (:345-348); `padding_mask = target_shifted != pad` (:353);
     `decoder(target_shifted, padding_mask, memory)`; CrossEntropyLoss(ignore_index=pad) against `tokens[:, 1:]`
     (:362-366; the `* padding_mask ... / padding_mask.sum()` that follows is the identity on a scalar);
     argmax decode for the metrics (:369-376).
  5. result dict of objectives.compute_vqa (objectives.py:121-147): vqa_loss, vqa_logits (= generated texts),
     vqa_labels (= target texts), vqa_scores, vqa_answer_types.  The CPU string metrics (ROUGE / BLEU / exact match)
     are outside the hot-path scope.

The HF `BertTokenizer('bert-base-uncased')` (:191) needs a vocabulary file that does not exist offline: pass any object
with the same surface (`__call__(texts, padding, truncation, return_tensors, max_length).input_ids`, `decode`,
`cls/sep/pad/eos_token_id`, `vocab_size`); `HashTokenizer` below is a self-contained stand-in with BERT's special ids.
"""
import torch
import torch.nn as nn

from . import functional as Fn
from .decoder import Decoder
from .m3ae_module import M3AETransformerSS, _Base


class HashTokenizer:
    """Whitespace tokenizer with BERT's special ids (pad 0, cls 101, sep 102) and a hashed vocabulary of `vocab_size`
    ids.  Stands in for BertTokenizer when the vocabulary file is unavailable (offline boxes, tests, benchmarks)."""

    cls_token_id, sep_token_id, pad_token_id, eos_token_id = 101, 102, 0, None

    def __init__(self, vocab_size=30522):
        self.vocab_size = vocab_size
        self._words = {}

    def _id(self, w):
        h = 1000 + (sum((i + 1) * ord(c) * 2654435761 for i, c in enumerate(w)) % (self.vocab_size - 1000))
        self._words.setdefault(h, w)
        return h

    def __call__(self, texts, padding=True, truncation=True, return_tensors="pt", max_length=12):
        rows = []
        for t in texts:
            ids = [self.cls_token_id] + [self._id(w) for w in t.lower().split()]
            ids = ids[:max_length - 1] + [self.sep_token_id]
            rows.append(ids)
        width = max(len(r) for r in rows)
        out = torch.full((len(rows), width), self.pad_token_id, dtype=torch.long)
        for i, r in enumerate(rows):
            out[i, :len(r)] = torch.tensor(r)
        return type("Enc", (), {"input_ids": out})()

    def decode(self, seq, skip_special_tokens=True):
        ids = seq.tolist() if torch.is_tensor(seq) else list(seq)
        special = {self.cls_token_id, self.sep_token_id, self.pad_token_id}
        return " ".join(self._words.get(i, f"[{i}]") for i in ids if not (skip_special_tokens and i in special))


class DecoderModel(_Base):
    def __init__(self, m3ae_config, max_answer_length=12, min_answer_length=1, freeze_m3ae=True, tokenizer=None,
                 compute_dtype=None, decoder_max_len=128):
        super().__init__()
        self.hparams = type("HParams", (), {})()
        self.hparams.m3ae_config = m3ae_config
        if tokenizer is None:
            try:
                from transformers import BertTokenizer
                tokenizer = BertTokenizer.from_pretrained("bert-base-uncased")  # m3ae_decoder.py:191
            except Exception:
                tokenizer = HashTokenizer()
        self.tokenizer = tokenizer
        self.m3ae = M3AETransformerSS(m3ae_config, compute_dtype)
        dt = self.m3ae.compute_dtype
        self.decoder = Decoder(num_layers=6, d_model=768, num_heads=8, d_ff=768 * 4, dropout=0.1, max_len=decoder_max_len,
                               target_vocab_size=self.tokenizer.vocab_size, compute_dtype=dt)  # :195-203
        path = m3ae_config.get("decoder_load_path", "")
        if path:  # :206-210
            self.decoder.load_state_dict(torch.load(path, map_location="cpu")["model_state_dict"], strict=True)
        if freeze_m3ae:  # :213-216
            for p in self.m3ae.parameters():
                p.requires_grad = False
        self.max_answer_length = max_answer_length
        self.min_answer_length = min_answer_length
        self.current_tasks = list()

    def memory(self, m3ae_output):
        """m3ae_decoder.py:302-313."""
        cfg = self.hparams.m3ae_config
        parts = []
        if cfg.get("mm_encoder_inputs_include_imagetext_feats", True):
            parts += [m3ae_output["multi_modal_image_feats"], m3ae_output["multi_modal_text_feats"]]
        if cfg.get("mm_encoder_inputs_include_cls_feats", True):
            parts.append(m3ae_output["multi_modal_cls_feats"].view(-1, 2, 768))
        return torch.cat(parts, dim=1)

    def forward(self, batch, test=False):
        m3ae_output = self.m3ae.infer(batch, mask_text=False, mask_image=False)
        memory = self.memory(m3ae_output)
        targets = [ans[0] for ans in batch["vqa_answer"]]
        tok = self.tokenizer
        ret = {}
        if len(self.current_tasks) == 0 or test:
            output = self.decoder.search_path(memory, tokenizer=tok)
            generated = [tok.decode(seq, skip_special_tokens=True) for seq in output.cpu()]
            loss = 0
            ret["generated_ids"] = output
        else:
            tokens = tok(targets, padding=True, truncation=True, return_tensors="pt", max_length=self.max_answer_length).input_ids
            tokens = tokens.to(memory.device)
            shifted = tokens[:, :-1]
            shifted = torch.where(shifted == tok.sep_token_id, torch.full_like(shifted, tok.pad_token_id), shifted)
            padding_mask = torch.ne(shifted, tok.pad_token_id)
            logits, _ = self.decoder(shifted, padding_mask, memory)
            golden = tokens[:, 1:]
            loss = Fn.cross_entropy(logits, golden.contiguous(), ignore_index=tok.pad_token_id)
            with torch.no_grad():
                generated = [tok.decode(seq, skip_special_tokens=True) for seq in logits.argmax(dim=-1).cpu()]
            ret["decoder_logits"] = logits
        phase = "test" if test else ("train" if self.training else "val")
        if torch.is_tensor(loss):
            self.log(f"{phase}/vqa/loss", loss.detach())
        ret.update({"vqa_loss": loss, "vqa_logits": [[g] for g in generated], "vqa_labels": targets,
                    "vqa_scores": batch.get("vqa_scores"),
                    "vqa_answer_types": torch.tensor(batch.get("answer_types", [0] * len(targets)), device=memory.device)})
        return ret

    # ---- pytorch_lightning hooks (m3ae_decoder.py:391-443)
    def _set_task(self):
        self.current_tasks = [k for k, v in self.hparams.m3ae_config["loss_names"].items() if v > 0]

    def training_step(self, batch, batch_idx):
        self._set_task()
        output = self(batch)
        total = sum(v * self.hparams.m3ae_config["loss_names"][k.replace("_loss", "")] for k, v in output.items() if "loss" in k)
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
        """m3ae_t5_utils.set_schedule_decoder (:290-380): two AdamW groups (decay / no decay), one learning rate,
        polynomial decay with warm-up."""
        c = self.hparams.m3ae_config
        lr, wd = c.get("learning_rate", 1e-5), c.get("weight_decay", 0.01)
        no_decay = ["bias", "LayerNorm.bias", "LayerNorm.weight", "norm.bias", "norm.weight"]
        named = [(n, p) for n, p in self.named_parameters()]
        groups = [{"params": [p for n, p in named if not any(nd in n for nd in no_decay)], "weight_decay": wd, "lr": lr},
                  {"params": [p for n, p in named if any(nd in n for nd in no_decay)], "weight_decay": 0.0, "lr": lr}]
        opt = torch.optim.AdamW(groups, lr=lr, eps=1e-8, betas=(0.9, 0.98))
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
