"""Model-level parity: the fusion section of M3AETransformerSS.infer + vqa_head + BCE loss (forward and backward)
on the CUDA path vs the oracle and the golden vectors from the unmodified reference; plus the drop-in surface
(batch dict -> infer outputs, pytorch_lightning-style fine-tune loop)."""
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import m3ae_b200  # noqa: E402,F401
from m3ae_b200.m3ae_module import M3AEFusion, M3AETransformerSS  # noqa: E402
from oracle import m3ae_oracle as O  # noqa: E402

DEV = "cuda"
GOLD = os.path.join(ROOT, "tests", "golden", "fusion.pt")
TOL = {torch.float32: 1e-4, torch.bfloat16: 2e-2}


def rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


def make_fusion(dtype, drop=0.0):
    m = M3AEFusion({"drop_rate": drop}, compute_dtype=dtype)
    missing, unexpected = m.load_state_dict(O.make_fusion_state(21), strict=True)
    return m.to(DEV)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_fusion_forward_logits_loss_argmax_match_reference_golden(dtype):
    gold = torch.load(GOLD)
    inp = O.make_fusion_inputs(22, B=2)
    m = make_fusion(dtype).eval()
    with torch.no_grad():
        loss, logits, ret = m.vqa_loss_from_feats(inp["text_feats"].to(DEV), inp["image_feats"].to(DEV),
                                                  inp["text_masks"].to(DEV), inp["targets"].to(DEV))
    tol = TOL[dtype]
    assert rel(ret["multi_modal_cls_feats"], gold["multi_modal_cls_feats"]) < tol
    assert rel(ret["multi_modal_text_feats"], gold["multi_modal_text_feats"]) < tol
    assert rel(ret["multi_modal_image_feats"][:, 0], gold["multi_modal_image_feats_tok0"]) < tol
    assert torch.equal(ret["extended_text_masks"].cpu(), gold["extended_text_masks"])
    assert rel(logits, gold["vqa_logits"]) < tol, f"logits {rel(logits, gold['vqa_logits']):.3e}"
    assert abs(loss.item() - gold["vqa_loss"].item()) / gold["vqa_loss"].item() < tol
    assert torch.equal(logits.argmax(1).cpu(), gold["vqa_logits"].argmax(1)), "argmax answers must be identical"


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_finetune_step_gradients_match_oracle(dtype):
    tol = 2e-4 if dtype == torch.float32 else 5e-2
    inp = O.make_fusion_inputs(23, B=2)
    sd = {k: v.clone().requires_grad_(True) for k, v in O.make_fusion_state(21).items()}
    to, io = inp["text_feats"].clone().requires_grad_(True), inp["image_feats"].clone().requires_grad_(True)
    loss_o, logits_o, _ = O.finetune_step(to, io, inp["text_masks"], inp["targets"], sd)
    loss_o.backward()
    m = make_fusion(dtype).eval()
    tc, ic = inp["text_feats"].to(DEV).requires_grad_(True), inp["image_feats"].to(DEV).requires_grad_(True)
    loss_c, logits_c, _ = m.vqa_loss_from_feats(tc, ic, inp["text_masks"].to(DEV), inp["targets"].to(DEV))
    loss_c.backward()
    assert abs(loss_c.item() - loss_o.item()) / abs(loss_o.item()) < TOL[dtype]
    assert rel(tc.grad, to.grad) < tol and rel(ic.grad, io.grad) < tol
    worst = ("", 0.0)
    for k, p in m.named_parameters():
        if k.endswith("self.key.bias"):
            continue
        assert p.grad is not None, k
        worst = max(worst, (k, rel(p.grad, sd[k].grad)), key=lambda t: t[1])
    assert worst[1] < tol, f"worst parameter gradient: {worst}"


def synthetic_batch(B, seed=1):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(3, 50265, (B, 32), generator=g)
    ids[:, 0] = 0
    lens = torch.randint(8, 33, (B,), generator=g)
    masks = (torch.arange(32)[None, :] < lens[:, None]).long()
    ids = torch.where(masks.bool(), ids, torch.ones_like(ids))
    labels = torch.randint(0, 498, (B,), generator=g)
    return {"image": [torch.randn(B, 3, 288, 288, generator=g).to(DEV)], "text_ids": ids.to(DEV), "text_masks": masks.to(DEV),
            "text_labels": torch.full((B, 32), -100).to(DEV), "vqa_labels": [[int(l)] for l in labels],
            "vqa_scores": [[1.0]] * B, "answer_types": [0] * B, "vqa_answer": [["yes"]] * B}


def test_m3ae_transformer_ss_batch_dict_to_infer_outputs_and_pl_loop():
    torch.manual_seed(0)
    model = M3AETransformerSS({"drop_rate": 0.1}, compute_dtype=torch.bfloat16).to(DEV)
    batch = synthetic_batch(2)
    model.eval()
    with torch.no_grad():
        ret = model.infer(batch)
    for k in ("images", "patched_images", "text_labels", "text_ids", "text_masks", "extended_image_masks",
              "extended_text_masks", "multi_modal_text_feats", "multi_modal_image_feats", "multi_modal_cls_feats"):
        assert k in ret, k
    assert ret["multi_modal_cls_feats"].shape == (2, 1536) and ret["multi_modal_image_feats"].shape == (2, 325, 768)
    assert ret["patched_images"].shape == (2, 324, 768) and ret["extended_image_masks"].shape == (2, 1, 1, 325)
    # pytorch_lightning 1.3.2-style manual loop: training_step -> backward -> optimizer/scheduler step (interval "step")
    model.train()
    (opt,), (sched,) = model.configure_optimizers(max_steps=10)
    names = {n for n, _ in model.named_parameters()}
    assert any(n.startswith("multi_modal_language_layers.0.attention.self.query") for n in names)
    w0 = model.multi_modal_vision_layers[0].output.dense.weight.detach().clone()
    losses = []
    for i in range(3):
        loss = model.training_step(batch, i)
        opt.zero_grad()
        loss.backward()
        opt.step()
        sched["scheduler"].step()
        losses.append(loss.item())
    assert all(torch.isfinite(torch.tensor(losses)))
    assert not torch.equal(w0, model.multi_modal_vision_layers[0].output.dense.weight.detach())
    out = model.test_step(batch, 0)
    assert out["vqa_logits"].shape == (2, 498)
