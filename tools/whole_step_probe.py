"""Host-issue time vs GPU time of the whole-model fine-tune step (CLIP ViT-B/16 + RoBERTa-base + fusion + heads), batch 64.
Usage (GPU box): python tools/whole_step_probe.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import m3ae_b200  # noqa: F401
from m3ae_b200 import ops
from m3ae_b200.m3ae_module import M3AETransformerSS
from oracle import m3ae_oracle as O

dev = torch.device("cuda")
B = int(os.environ.get("B", "64"))
torch.manual_seed(0)
model = M3AETransformerSS({"drop_rate": 0.1}, compute_dtype=torch.bfloat16).to(dev).train()
opt = torch.optim.AdamW(list(model.parameters()), lr=1e-5, betas=(0.9, 0.98), eps=1e-8, weight_decay=0.01, fused=True)
batches = [O.make_model_batch(70 + i, B, device=dev) for i in range(2)]


def step(i):
    opt.zero_grad(set_to_none=True)
    loss = model.training_step(batches[i % 2], i)
    loss.backward()
    opt.step()


for i in range(3):
    step(i)
torch.cuda.synchronize()
n0 = ops.launch_count()
host = []
for i in range(3):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    step(i)
    host.append((time.perf_counter() - t0) * 1e3)
    torch.cuda.synchronize()
print("library launches per step", (ops.launch_count() - n0) / 3, "host issue ms per step", [round(h, 1) for h in host])
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
for i in range(5):
    step(i)
e1.record()
torch.cuda.synchronize()
print("gpu ms per step", round(e0.elapsed_time(e1) / 5, 2))
if os.environ.get("PROFILE", "0") == "1":
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        step(0)
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=40, max_name_column_width=70))

# ---- host-issue breakdown (no synchronisation inside a step: pure Python / launch overhead per segment)
import collections
seg = collections.defaultdict(float)
torch.cuda.synchronize()
for i in range(3):
    b = batches[i % 2]
    t = time.perf_counter(); opt.zero_grad(set_to_none=True); seg["zero_grad"] += time.perf_counter() - t
    t = time.perf_counter(); ut = model.encode_text(b["text_ids"], b["text_masks"]); seg["text encoder fwd"] += time.perf_counter() - t
    t = time.perf_counter(); ui = model.vision_encoder(b["image"][0]); seg["vision encoder fwd"] += time.perf_counter() - t
    t = time.perf_counter(); ret = model.fuse(ut, ui, b["text_masks"], 1, False); seg["fusion fwd"] += time.perf_counter() - t
    t = time.perf_counter()
    logits = model.vqa_head(ret["multi_modal_cls_feats"])
    targets = torch.zeros(len(logits), 498, device=dev)
    from m3ae_b200 import functional as Fn
    loss = Fn.vqa_bce_loss(logits, targets)
    seg["head + loss"] += time.perf_counter() - t
    t = time.perf_counter(); loss.backward(); seg["backward (all)"] += time.perf_counter() - t
    t = time.perf_counter(); opt.step(); seg["optimizer step"] += time.perf_counter() - t
    torch.cuda.synchronize()
print("host ms per step by segment:", {k: round(v / 3 * 1e3, 2) for k, v in seg.items()})
