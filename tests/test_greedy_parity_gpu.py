"""north_star: "identical greedy-decoded token IDs on at least 99.9% of samples" for the bf16 product path (VERDICT r1
'weak' #1).

Random-init weights give near-tied vocabulary logits (the top-2 gap of 30 522 i.i.d. logits is routinely below bf16
resolution), which says nothing about a fine-tuned head.  So the fixture here IS a fine-tuned head: the decoder is trained
on the GPU for a few hundred steps (through the drop-in's own bf16 training path, dropout on) on a synthetic answer-
generation task — 32 answer sequences, each tied to a memory prototype — until it decodes them; then 1 024 fresh noisy
memories (8 batches of 128 x 359 tokens, BASELINE config 3's shape) are decoded by
  (a) the CPU oracle (reference algorithm, m3ae_decoder.py:141-182) on a subset,
  (b) the fp32 CUDA path, (c) the bf16 CUDA path,
and (b) must equal (a) exactly, (c) must equal (b) on >= 99.9 % of the sequences.  A second test keeps the random-init case
honest: every bf16/fp32 divergence there must be a near-tie of the fp32 logits (gap below the 2e-2 bf16 gate)."""
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import m3ae_b200  # noqa: E402,F401
from m3ae_b200.decoder import Decoder  # noqa: E402
from oracle import m3ae_oracle as O  # noqa: E402

DEV = "cuda"
LM, L = 359, 16


class Tok:
    cls_token_id, sep_token_id, eos_token_id, pad_token_id = 101, 102, None, 0


def make(sd, dtype):
    d = Decoder(6, 768, 8, 3072, 0.1, L, 30522, compute_dtype=dtype)
    d.load_state_dict(sd, strict=True)
    return d.to(DEV)


def task(n_cls=32, seed=7):
    g = torch.Generator().manual_seed(seed)
    protos = torch.randn(n_cls, LM, 768, generator=g)
    lens = torch.randint(2, 9, (n_cls,), generator=g)
    answers = torch.zeros(n_cls, 10, dtype=torch.long)
    for k in range(n_cls):
        n = int(lens[k])
        answers[k, 0] = 101
        answers[k, 1:1 + n] = torch.randint(1000, 30000, (n,), generator=g)
        answers[k, 1 + n] = 102
    return protos, answers


def batch(protos, answers, B, gen, noise=0.5):
    cls = torch.randint(0, protos.shape[0], (B,), generator=gen)
    mem = protos[cls] + noise * torch.randn(B, LM, 768, generator=gen)
    return mem, answers[cls], cls


def finetune(dec, protos, answers, steps=300):
    dec.train()
    opt = torch.optim.Adam([p for n, p in dec.named_parameters() if not (n.startswith("dec_layers.") and not n.startswith("dec_layers.5."))], lr=5e-4)
    gen = torch.Generator().manual_seed(11)
    loss = None
    for i in range(steps):
        mem, tok, _ = batch(protos, answers, 64, gen)
        tok = tok.to(DEV)
        shifted = tok[:, :-1]
        shifted = torch.where(shifted == 102, torch.zeros_like(shifted), shifted)
        logits, _ = dec(shifted, shifted != 0, mem.to(DEV))
        loss = torch.nn.functional.cross_entropy(logits.reshape(-1, logits.shape[-1]), tok[:, 1:].reshape(-1), ignore_index=0)
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
    return loss.item()


def test_finetuned_head_bf16_greedy_ids_identical_on_999_permille_of_1024_samples():
    protos, answers = task()
    dec = make(O.make_decoder_state(41), torch.bfloat16)
    final_loss = finetune(dec, protos, answers)
    sd = {k: v.detach().float().cpu() for k, v in dec.state_dict().items()}
    d32, d16 = make(sd, torch.float32).eval(), make(sd, torch.bfloat16).eval()
    gen = torch.Generator().manual_seed(99)
    same = solved = total = 0
    first = True
    for _ in range(8):
        mem, tok, _ = batch(protos, answers, 128, gen)
        ids32 = d32.search_path(mem.to(DEV), Tok).cpu()
        ids16 = d16.search_path(mem.to(DEV), Tok).cpu()
        want = torch.zeros(128, L, dtype=torch.long)
        want[:, :9] = tok[:, 1:]
        same += int((ids32 == ids16).all(dim=1).sum())
        solved += int((ids32 == want).all(dim=1).sum())
        total += 128
        if first:  # the reference algorithm itself (CPU oracle, whole-prefix re-run, no cache) on a subset
            first = False
            sdo = {k: v for k, v in sd.items() if not (k.startswith("dec_layers.") and not k.startswith("dec_layers.5."))}
            with torch.no_grad():
                ref = O.search_path(mem[:16], sdo, L)
            assert torch.equal(ids32[:16], ref), "fp32 CUDA greedy ids must equal the oracle's"
    print(f"fine-tuned head: final loss {final_loss:.4f}; fp32 decodes the target answer on {solved}/{total}; "
          f"bf16 == fp32 sequences on {same}/{total} ({100.0 * same / total:.2f} %)")
    assert solved >= 0.95 * total, "the fixture did not train: margins would not be those of a fine-tuned head"
    assert same >= 0.999 * total, f"bf16 greedy ids identical on only {same}/{total} sequences (north_star: >= 99.9 %)"


def test_random_init_divergences_are_fp32_near_ties():
    """Random-init weights: wherever the bf16 sequence first departs from the fp32 one, the token bf16 chose must be a
    near-tie in the fp32 logits at that step (within the 2e-2 bf16 gate, relative to the logits' spread)."""
    sd = O.make_decoder_state(41)
    d32, d16 = make(sd, torch.float32).eval(), make(sd, torch.bfloat16).eval()
    mem = torch.randn(128, LM, 768, generator=torch.Generator().manual_seed(5)).to(DEV)
    ids32, ids16 = d32.search_path(mem, Tok), d16.search_path(mem, Tok)
    differ = (ids32 != ids16).any(dim=1).nonzero().flatten().tolist()
    identical = 128 - len(differ)
    worst = 0.0
    for b in differ:
        t = int((ids32[b] != ids16[b]).nonzero()[0])
        prefix = torch.cat([torch.tensor([101], device=DEV), ids32[b, :t]])[None]
        with torch.no_grad():
            logits, _ = d32(prefix, None, mem[b:b + 1])
        row = logits[0, -1]
        gap = (row.max() - row[ids16[b, t]]).item() / (row.max() - row.mean()).item()
        worst = max(worst, gap)
    print(f"random-init: {identical}/128 identical sequences; worst relative fp32 gap at a divergence {worst:.4f}")
    assert worst < 2e-2, "a bf16 divergence is not explained by an fp32 near-tie"
