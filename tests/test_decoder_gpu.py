"""Generative decoder head on the CUDA path vs the oracle restatement of the reference Decoder and vs golden vectors
from the unmodified reference (tests/golden/decoder.pt)."""
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
from oracle import make_golden as G  # noqa: E402

DEV = "cuda"
GOLD = os.path.join(ROOT, "tests", "golden", "decoder.pt")


def rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


def make_decoder(sd, dtype, max_len=16):
    d = Decoder(6, 768, 8, 3072, 0.1, max_len, 30522, compute_dtype=dtype)
    d.load_state_dict(sd, strict=True)
    return d.to(DEV).eval()


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-4), (torch.bfloat16, 2e-2)])
def test_teacher_forced_logits_match_oracle_and_golden(dtype, tol):
    gold = torch.load(GOLD)
    sd = O.make_decoder_state(31)
    mem, ids, pm = G.decoder_inputs(32, B=4, T=9)
    with torch.no_grad():
        ref, att_ref = O.decoder_forward(ids, pm, mem, sd)
    dec = make_decoder(sd, dtype)
    logits, att = dec(ids.to(DEV), pm.to(DEV), mem.to(DEV))
    assert logits.shape == ref.shape and logits.dtype == torch.float32
    assert rel(logits, ref) < tol, f"logits: {rel(logits, ref):.3e}"
    assert rel(logits[:, :, ::16], gold["logits_sub"]) < tol
    assert rel(att["layer6_dec_self"], att_ref["layer6_dec_self"]) < tol
    assert rel(att["layer6_dec_cross"], att_ref["layer6_dec_cross"]) < tol
    if dtype == torch.float32:
        assert torch.equal(logits.argmax(-1).cpu(), gold["argmax"])


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_teacher_forced_backward_matches_oracle(dtype):
    # bf16: weight gradients here are sums over only 21 token rows of bf16 activations x bf16 upstream grads (plus ReLU-mask
    # flips where the pre-activation is within bf16 rounding of 0): gate them at 1e-1; the fp32 path is held to 1e-4.
    tol = 1e-4 if dtype == torch.float32 else 1e-1
    sd = O.make_decoder_state(31)
    mem, ids, pm = G.decoder_inputs(33, B=3, T=7)
    tgt = torch.randint(0, 30522, (3, 7), generator=torch.Generator().manual_seed(1))
    sdo = {k: v.clone().requires_grad_(v.dtype.is_floating_point and k != "positional_encoding.pe") for k, v in sd.items()}
    mo = mem.clone().requires_grad_(True)
    lo, _ = O.decoder_forward(ids, pm, mo, sdo)
    torch.nn.functional.cross_entropy(lo.reshape(-1, lo.shape[-1]), tgt.reshape(-1)).backward()
    dec = make_decoder(sd, dtype)  # eval mode: dropout off (train mode applies the reference's five dropouts), grads flow
    mc = mem.to(DEV).requires_grad_(True)
    lc, _ = dec(ids.to(DEV), pm.to(DEV), mc)
    torch.nn.functional.cross_entropy(lc.reshape(-1, lc.shape[-1]), tgt.to(DEV).reshape(-1)).backward()
    assert rel(mc.grad, mo.grad) < tol, f"d memory: {rel(mc.grad, mo.grad):.3e}"
    live = "dec_layers.5."
    for k, p in dec.named_parameters():
        if k.startswith("dec_layers.") and not k.startswith(live):
            assert p.grad is None, f"{k}: dead layers must receive no gradient"
            continue
        if k.endswith("mha1.in_proj_bias") or k.endswith("mha2.in_proj_bias"):
            # the key-bias third of the packed bias has an identically zero gradient: compare q and v parts
            g, go = p.grad.view(3, -1), sdo[k].grad.view(3, -1)
            assert rel(g[0], go[0]) < tol and rel(g[2], go[2]) < tol, k
            continue
        assert rel(p.grad, sdo[k].grad) < tol, f"{k}: {rel(p.grad, sdo[k].grad):.3e}"


class Tok:
    cls_token_id, sep_token_id, eos_token_id, pad_token_id = 101, 102, None, 0


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_greedy_kv_cached_ids_match_reference_golden(dtype):
    gold = torch.load(GOLD)
    sd = O.make_decoder_state(31)
    mem, _, _ = G.decoder_inputs(32, B=4, T=9)
    dec = make_decoder(sd, dtype)
    ids = dec.search_path(mem.to(DEV), Tok).cpu()
    assert ids.shape == (4, 16) and ids.dtype == torch.int64
    if dtype == torch.float32:
        assert torch.equal(ids, gold["greedy"])
    else:  # bf16: a near-tie may flip one argmax; once a token differs the suffix diverges, so compare prefixes
        assert (ids[:, 0] == gold["greedy"][:, 0]).all()
    sd2 = dict(sd)
    sd2["final_linear.bias"] = sd["final_linear.bias"].clone()
    sd2["final_linear.bias"][102] += 4.0
    dec2 = make_decoder(sd2, dtype)
    ids2 = dec2.search_path(mem.to(DEV), Tok).cpu()
    ids3 = dec2.search_path(mem.to(DEV), Tok, early_exit_every=2).cpu()
    if dtype == torch.float32:
        assert torch.equal(ids2, gold["greedy_sep"])
    assert torch.equal(ids2, ids3), "the all-finished early exit must not change the result"


def test_greedy_batch128_matches_oracle_fp32_and_bf16_agreement():
    """BASELINE config 3 shape: B=128 memories of 359 tokens, up to 16 tokens.  fp32 CUDA path must reproduce the oracle's
    ids exactly; the bf16 path is reported as per-sample agreement (north_star: >= 99.9% identical greedy IDs is
    the target for trained weights; with random-init weights logit gaps are tiny, so the bf16 figure is informational)."""
    sd = O.make_decoder_state(41, live_only=True)
    full = O.make_decoder_state(41)
    g = torch.Generator().manual_seed(5)
    mem = torch.randn(128, 359, 768, generator=g)
    with torch.no_grad():
        ref = O.search_path(mem[:16], sd, 16)
    dec = make_decoder(full, torch.float32)
    ids = dec.search_path(mem.to(DEV), Tok).cpu()
    assert torch.equal(ids[:16], ref)
    dec16 = make_decoder(full, torch.bfloat16)
    ids16 = dec16.search_path(mem.to(DEV), Tok).cpu()
    agree = (ids16 == ids).all(dim=1).float().mean().item()
    first = (ids16[:, 0] == ids[:, 0]).float().mean().item()
    print(f"bf16 vs fp32 greedy: identical sequences {agree:.3f}, identical first token {first:.3f}")
    assert first > 0.9


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_greedy_cuda_graph_replay_equals_eager_program(dtype):
    """search_path replays a captured CUDA graph of the greedy program: it must return exactly what the eager program
    returns — on the capturing call, on replays with other inputs, after an in-place change of an fp32 master
    (read at replay time) and after a differentiated forward (operand cache invalidated -> re-capture)."""
    sd = O.make_decoder_state(31)
    dec = make_decoder(sd, dtype)
    g = torch.Generator().manual_seed(77)
    mems = [torch.randn(4, 359, 768, generator=g).to(DEV) for _ in range(3)]
    for m in mems + [mems[0]]:  # first call captures, the rest replay
        assert torch.equal(dec.search_path(m, Tok), dec.search_path(m, Tok, use_graph=False))
    assert len(dec._graphs) == 1
    with torch.no_grad():
        dec.final_linear.bias[102] += 6.0  # makes [SEP] win early: the output must change and still match eager
    a, b = dec.search_path(mems[1], Tok), dec.search_path(mems[1], Tok, use_graph=False)
    assert torch.equal(a, b) and (a == 102).any()
    assert len(dec._graphs) == 1
    # a training forward invalidates the bf16 operand cache (the optimizer will move the weights): graphs are dropped
    tgt = torch.randint(1000, 30000, (4, 6), device=DEV)
    dec(tgt, None, mems[0])[0].float().square().mean().backward()
    with torch.no_grad():
        dec.dec_layers[-1].ffn[0].weight.mul_(1.5)  # stands in for an optimizer step
    dec.eval()
    assert len(dec._graphs) == 0
    assert torch.equal(dec.search_path(mems[2], Tok), dec.search_path(mems[2], Tok, use_graph=False))


# ----------------------------------------------------------------------------- train-mode dropout (m3ae_decoder.py:42-43,62,80,86,128)
def test_train_mode_applies_the_reference_dropouts():
    """The reference fine-tunes the decoder with p = 0.1 dropout on the embedding, inside both MultiheadAttentions and
    after each sub-block.  Philox masks cannot be bit-matched to torch's; checked here: train != eval, two train forwards
    differ, the same torch seed reproduces the forward exactly, dropout's effect has the right size (p -> 0 limit and
    variance growth), gradients are finite and reach every live parameter, eval mode is untouched."""
    sd = O.make_decoder_state(31)
    mem, ids, pm = G.decoder_inputs(35, B=8, T=9)
    mem, ids, pm = mem.to(DEV), ids.to(DEV), pm.to(DEV)
    dec = make_decoder(sd, torch.bfloat16)
    with torch.no_grad():
        ev, _ = dec(ids, pm, mem)
        ev2, _ = dec(ids, pm, mem)
    assert torch.equal(ev, ev2)
    dec.train()
    torch.manual_seed(123)
    with torch.no_grad():
        t1, _ = dec(ids, pm, mem)
        t2, _ = dec(ids, pm, mem)
    import m3ae_b200.functional as Fn
    import itertools
    torch.manual_seed(123)
    Fn._seed_counter = itertools.count(next(Fn._seed_counter) - 2)   # rewind the per-call seed counter by the two calls above
    with torch.no_grad():
        t1b, _ = dec(ids, pm, mem)
    assert torch.equal(t1, t1b), "same torch seed + same call index must give the same dropout masks"
    d_te, d_tt = rel(t1, ev), rel(t1, t2)
    print(f"decoder dropout: train vs eval {d_te:.3f}, train vs train {d_tt:.3f}")
    assert 0.02 < d_te < 0.8 and 0.02 < d_tt < 1.0
    # expectation over masks approaches the eval output: average of many train forwards is closer to eval than one is
    with torch.no_grad():
        avg = torch.stack([dec(ids, pm, mem)[0] for _ in range(24)]).mean(0)
    assert rel(avg, ev) < 0.6 * d_te
    dec.dropout.p = 0.0  # p = 0 everywhere -> identical to eval
    for l in dec.dec_layers:
        for m_ in (l.dropout1, l.dropout2, l.dropout3):
            m_.p = 0.0
    # the attention dropout is read from self.dropout.p as well (one `dropout` constructor argument feeds all sites)
    with torch.no_grad():
        t0, _ = dec(ids, pm, mem)
    assert torch.equal(t0, ev)
    dec.dropout.p = 0.1
    logits, _ = dec(ids, pm, mem.requires_grad_(True))
    torch.nn.functional.cross_entropy(logits.reshape(-1, logits.shape[-1]), ids.reshape(-1)).backward()
    assert torch.isfinite(mem.grad).all() and mem.grad.abs().sum() > 0
    for k, p in dec.named_parameters():
        if k.startswith("dec_layers.5.") or k.startswith("final_linear") or k.startswith("target_embedding"):
            assert p.grad is not None and torch.isfinite(p.grad).all(), k


def test_linear_epilogue_dropout_backward_uses_the_forward_mask():
    """y = drop(x W^T + b) + r through functional.linear: dL/dx, dL/dW, dL/db must be those of the SAME mask (recovered
    from the forward output) — checked against torch autograd with that mask plugged in."""
    import m3ae_b200.functional as Fn
    g = torch.Generator().manual_seed(5)
    M, K, N, p = 96, 768, 768, 0.25
    x = torch.randn(M, K, generator=g).to(DEV).requires_grad_(True)
    w = (torch.randn(N, K, generator=g) * 0.05).to(DEV).requires_grad_(True)
    b = (torch.randn(N, generator=g) * 0.1).to(DEV).requires_grad_(True)
    r = torch.randn(M, N, generator=g).to(DEV).requires_grad_(True)
    y = Fn.linear(x, w, b, dtype=torch.float32, residual=r, drop_p=p, drop_seed=77, drop_site=3)
    with torch.no_grad():
        pre = x @ w.t() + b
        mask = ((y - r).abs() > 1e-12).float()
        keep = mask.mean().item()
        scale = ((y - r)[mask.bool()] / pre[mask.bool()]).median().item()
    assert abs(keep - (1 - p)) < 0.02 and abs(scale - 1 / (1 - p)) < 2e-3
    probe = torch.randn(M, N, generator=torch.Generator().manual_seed(6)).to(DEV)
    (y * probe).sum().backward()
    xr, wr, br, rr = (t.detach().clone().requires_grad_(True) for t in (x, w, b, r))
    yr = (xr @ wr.t() + br) * mask * scale + rr
    (yr * probe).sum().backward()
    for a, c, n in ((x, xr, "dx"), (w, wr, "dw"), (b, br, "db"), (r, rr, "dres")):
        assert rel(a.grad, c.grad) < 1e-4, (n, rel(a.grad, c.grad))


def test_attention_maps_are_lazy_and_cover_all_layers():
    sd = O.make_decoder_state(31)
    mem, ids, pm = G.decoder_inputs(32, B=4, T=9)
    with torch.no_grad():
        _, att_ref = O.decoder_forward(ids, pm, mem, sd, all_layers=True)
    dec = make_decoder(sd, torch.float32)
    from m3ae_b200 import ops
    n0 = ops.launch_count()
    with torch.no_grad():
        logits, att = dec(ids.to(DEV), pm.to(DEV), mem.to(DEV))
    n_fwd = ops.launch_count() - n0
    assert sorted(att.keys()) == sorted(att_ref.keys()) and len(att) == 12
    n0 = ops.launch_count()
    a6 = att["layer6_dec_cross"]
    assert 0 < ops.launch_count() - n0 <= 2 and n_fwd < 40
    assert rel(a6, att_ref["layer6_dec_cross"]) < 1e-4
    assert rel(att["layer2_dec_self"], att_ref["layer2_dec_self"]) < 1e-4     # dead layer: evaluated on demand
    assert rel(att["layer2_dec_cross"], att_ref["layer2_dec_cross"]) < 1e-4


def test_search_path_sees_an_optimizer_step_that_follows_in_step_decoding():
    """forward -> backward -> search_path (in-step metrics) -> opt.step() -> search_path must decode the
    second time with the UPDATED weights (fused AdamW does not bump parameter versions)."""
    sd = O.make_decoder_state(31)
    dec = make_decoder(sd, torch.bfloat16).train()
    opt = torch.optim.AdamW(dec.parameters(), lr=5e-2, fused=True)
    mem, ids, pm = G.decoder_inputs(36, B=4, T=9)
    mem, ids, pm = mem.to(DEV), ids.to(DEV), pm.to(DEV)
    logits, _ = dec(ids, pm, mem)
    torch.nn.functional.cross_entropy(logits.reshape(-1, logits.shape[-1]), ids.reshape(-1)).backward()
    dec.eval()
    before = dec.search_path(mem, Tok)
    opt.step()
    after = dec.search_path(mem, Tok)
    fresh = make_decoder({k: v.detach().cpu() for k, v in dec.state_dict().items()}, torch.bfloat16)
    assert torch.equal(after, fresh.search_path(mem, Tok)), "search_path decoded with stale operands after opt.step()"
    assert not torch.equal(before, after)
