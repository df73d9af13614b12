"""Layer-level parity of the CUDA path (drop-in BertCrossLayer over the C-ABI kernels) against the oracle
(oracle/m3ae_oracle.py, CPU fp32) and against golden vectors produced by the unmodified reference.

Gates (BASELINE.json north_star): fp32 path within 1e-4 relative, bf16 path within 2e-2 relative.
Relative error = ||a-b||_2 / ||b||_2 over the whole tensor.
"""
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import m3ae_b200  # noqa: E402,F401
from m3ae_b200.bert_cross_layer import BertCrossLayer, fused_layer_pair  # noqa: E402
from oracle import m3ae_oracle as O  # noqa: E402
from oracle import make_golden as G  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
DEV = "cuda"
TOL = {torch.float32: 1e-4, torch.bfloat16: 2e-2}


def rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


def fusion_cfg(drop=0.0):
    from transformers import RobertaConfig
    return RobertaConfig(vocab_size=50265, hidden_size=768, num_hidden_layers=6, num_attention_heads=12,
                         intermediate_size=3072, max_position_embeddings=32, hidden_dropout_prob=drop,
                         attention_probs_dropout_prob=drop)


def make_layer(seed, dtype, drop=0.0):
    layer = BertCrossLayer(fusion_cfg(drop), compute_dtype=dtype)
    layer.load_state_dict(O.make_cross_layer_state(seed), strict=True)
    return layer.to(DEV)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("case", G.cross_layer_cases(), ids=lambda c: c[0])
def test_cross_layer_matches_oracle_and_reference_golden(case, dtype):
    name, seed, B, Lq, Lk, qm, km = case
    tol = TOL[dtype]
    gold = torch.load(os.path.join(GOLD, "cross_layer.pt"))[name]
    x, y, mq, mk = G.cross_layer_inputs(seed, B, Lq, Lk, qm, km)
    emq, emk = O.extended_attention_mask(mq), O.extended_attention_mask(mk)
    # oracle (CPU fp32) forward + backward
    sd = {k: v.clone().requires_grad_(True) for k, v in O.make_cross_layer_state(seed).items()}
    xo, yo = x.clone().requires_grad_(True), y.clone().requires_grad_(True)
    out_o, sp_o, cp_o = O.bert_cross_layer(xo, yo, sd, "", emq, emk)
    probe = G.probe(out_o.shape, seed + 100)
    (out_o * probe).sum().backward()
    # CUDA path
    layer = make_layer(seed, dtype).eval()
    xc, yc = x.to(DEV).requires_grad_(True), y.to(DEV).requires_grad_(True)
    out_c, sp_c, cp_c = layer(xc, yc, emq.to(DEV), emk.to(DEV), output_attentions=True)
    assert out_c.dtype == torch.float32 and out_c.shape == out_o.shape
    (out_c * probe.to(DEV)).sum().backward()
    s = gold["tok_stride"]
    assert rel(out_c, out_o) < tol, f"layer out vs oracle: {rel(out_c, out_o):.3e}"
    assert rel(out_c[:, ::s], gold["layer_out"]) < tol, "layer out vs reference golden"
    assert rel(sp_c, sp_o) < tol and rel(cp_c, cp_o) < tol
    assert rel(xc.grad, xo.grad) < 2 * tol, f"dx: {rel(xc.grad, xo.grad):.3e}"
    assert rel(yc.grad, yo.grad) < 2 * tol, f"dy: {rel(yc.grad, yo.grad):.3e}"
    assert rel(xc.grad[:, ::s], gold["dx"]) < 2 * tol
    worst = ("", 0.0)
    for k, p in layer.named_parameters():
        if k.endswith("self.key.bias"):
            continue  # identically zero gradient (softmax shift invariance): rounding noise on both sides
        e = rel(p.grad, sd[k].grad)
        worst = max(worst, (k, e), key=lambda t: t[1])
    assert worst[1] < 2 * tol, f"worst param grad {worst}"


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_fused_text_image_pair_equals_separate_layers(dtype):
    B = 3
    g = torch.Generator().manual_seed(7)
    x = torch.randn(B, 32, 768, generator=g).to(DEV)
    y = torch.randn(B, 325, 768, generator=g).to(DEV)
    tm = torch.ones(B, 32, dtype=torch.long)
    tm[0, 9:] = 0
    emx = O.extended_attention_mask(tm).to(DEV)
    emy = O.extended_attention_mask(torch.ones(B, 325, dtype=torch.long)).to(DEV)
    lt, li = make_layer(1, dtype).eval(), make_layer(2, dtype).eval()
    xs, ys = x.clone().requires_grad_(True), y.clone().requires_grad_(True)
    x1 = lt(xs, ys, emx, emy)[0]
    y1 = li(ys, xs, emy, emx)[0]
    (x1.sum() + (y1 * y1).sum()).backward()
    named = [(k, p) for l in (lt, li) for k, p in l.named_parameters()]
    gs = [p.grad.clone() for _, p in named]
    for _, p in named:
        p.grad = None
    xp, yp = x.clone().requires_grad_(True), y.clone().requires_grad_(True)
    (x2,), (y2,) = fused_layer_pair(lt, li, xp, yp, emx, emy)
    (x2.sum() + (y2 * y2).sum()).backward()
    assert torch.equal(x1, x2) and torch.equal(y1, y2)
    assert rel(xp.grad, xs.grad) < 1e-5 and rel(yp.grad, ys.grad) < 1e-5
    for (k, p), gref in zip(named, gs):
        if k.endswith("self.key.bias"):
            continue  # identically-zero gradient (softmax shift invariance): rounding noise only
        assert rel(p.grad, gref) < 1e-3, k  # split-K atomics reorder fp32 sums


def test_training_mode_dropout_is_consistent():
    """Dropout cannot match PyTorch's RNG stream bit-for-bit; check that train mode perturbs the output, keeps the
    scale (E[out] ~ eval out) and that backward runs with the regenerated masks (finite, non-zero grads)."""
    layer = make_layer(3, torch.bfloat16, drop=0.1).train()
    g = torch.Generator().manual_seed(9)
    x = torch.randn(4, 32, 768, generator=g).to(DEV).requires_grad_(True)
    y = torch.randn(4, 325, 768, generator=g).to(DEV).requires_grad_(True)
    o1 = layer(x, y)[0]
    o2 = layer(x, y)[0]
    assert not torch.equal(o1, o2), "two training calls must draw different dropout masks"
    oe = layer.eval()(x, y)[0]
    assert rel(o1, oe) < 0.5 and rel(o1, oe) > 1e-3
    layer.train()
    layer(x, y)[0].square().mean().backward()
    for k, p in layer.named_parameters():
        assert torch.isfinite(p.grad).all(), k
    assert x.grad.abs().sum() > 0 and y.grad.abs().sum() > 0


def test_cpu_tensors_fail_loudly():
    layer = BertCrossLayer(fusion_cfg())
    with pytest.raises(RuntimeError, match="no CPU"):
        layer(torch.zeros(1, 32, 768), torch.zeros(1, 325, 768))


@pytest.mark.parametrize("fused", [True, False])
def test_operand_cache_follows_optimizer_steps(fused):
    """The bf16 operand cache must track the fp32 masters through optimizer steps.  torch's fused AdamW updates
    parameters WITHOUT bumping their version counters, so version-keyed caching would silently keep training (and then
    evaluating) on the initial weights."""
    layer = make_layer(5, torch.bfloat16).train()
    opt = torch.optim.AdamW(layer.parameters(), lr=1e-2, fused=fused)
    g = torch.Generator().manual_seed(11)
    x = torch.randn(2, 32, 768, generator=g).to(DEV)
    y = torch.randn(2, 325, 768, generator=g).to(DEV)
    for _ in range(2):
        opt.zero_grad(set_to_none=True)
        layer(x, y)[0].float().square().mean().backward()
        opt.step()
    fresh = BertCrossLayer(fusion_cfg(), compute_dtype=torch.bfloat16).to(DEV)
    fresh.load_state_dict(layer.state_dict())
    with torch.no_grad():
        o_eval = layer.eval()(x, y)[0]          # first no-grad forward after training: must re-cast
        o_eval2 = layer(x, y)[0]                # second: may use the cache
        o_fresh = fresh.eval()(x, y)[0]
    assert torch.equal(o_eval, o_fresh) and torch.equal(o_eval2, o_fresh)
    o_train = layer.train()(x.requires_grad_(True), y)[0]
    assert torch.equal(o_train.detach(), o_fresh)
    sd0 = O.make_cross_layer_state(5)
    assert rel(layer.intermediate.dense.weight, sd0["intermediate.dense.weight"]) > 1e-3  # the masters did move
