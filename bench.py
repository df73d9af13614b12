#!/usr/bin/env python
"""Benchmark of the M3AE fusion fine-tune hot path on B200 (BASELINE.json config[1]).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one fine-tune pass of the hot path over one synthetic batch of 64 samples per GPU:
modality projections -> 6+6 co-attention layers (fwd) -> poolers -> vqa_head -> BCE loss -> backward through all of
it -> (N>1: NCCL gradient all-reduce) -> AdamW step.  Dropout (p=0.1) is active as in the reference's fine-tune.
`value` is samples/s with inputs resident in HBM; `e2e` re-measures the same step driven from pinned HOST buffers
(H2D of every step's inputs — prefetched on a copy stream while the previous step runs — + D2H of the loss inside the
timed region).  `roofline` is the tcgen05 GEMM kernel (dominant kernel): algorithmic FLOPs of every GEMM launch in a
step / summed CUDA-event durations, against the measured bf16 peak in MEASURED_PEAKS.json.
`cpu_baseline` / `--impl reference` time the REFERENCE'S OWN modules (oracle/_ref: a git-ignored snapshot of
<reference>/m3ae/modules made by oracle/snapshot_ref.py, imported through oracle/ref_shim.py; kind "reference") on the
box's host cores — M3AETransformerSS.infer's fusion section + vqa_head + BCE + backward + AdamW on a bounded sample —
or, when the snapshot is absent, the oracle port (kind "port").  Both legs use ONE definition of the CPU step.

At N = 1 the same JSON line also carries:
  "decoder"                BASELINE config 3: greedy decode of 16 tokens, batch 128 (roofline vs HBM, e2e, cpu_baseline =
                           the reference Decoder.search_path at batch 16)
  "whole_step"             config 2 with the uni-modal encoders included (CLIP ViT-B/16 + RoBERTa-base on the same kernels)
  "classifier_forward_b8"  config 1: infer + vqa_head at batch 8, next to the reference's CPU fp32 time
  "t5_head"                config 4: T5 generative head, teacher-forced fine-tune step at batch 64 (frozen M3AE on the kernels, HF T5-small)
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

# The ONE JSON line goes to the real stdout; everything else a run prints (the reference's own modules print while they are
# constructed, m3ae_module.py:133) is sent to stderr so that the driver parses a clean line.
# (file-descriptor level: NCCL prints its version banner from C)
sys.stdout.flush()
_STDOUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)
sys.stdout = sys.stderr

LT, LI, HID, NLAB = 32, 325, 768, 498  # 32 text tokens, 325 image tokens (288 px / 16 + CLS), hidden 768, 498 answers
FLOPS_FWD_PER_SAMPLE = 42.778e9  # fusion stack forward, algorithmic 2*M*N*K (+attention)
FLOPS_FWD_BWD_PER_SAMPLE = 128.33e9


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"tflops": d.get("bf16_tflops_sustained", d.get("bf16_tflops")), "tflops_burst": d.get("bf16_tflops"),
                "hbm_gbs": d.get("hbm_gbs"), "source": "measured (MEASURED_PEAKS.json, sustained bf16)"}
    return {"tflops": 1400.0, "tflops_burst": 1590.0, "hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}


def gemm_traffic():
    """Per-launch DRAM bytes of the dominant kernel from the committed `ncu --set full` capture (tools/gemm_traffic.py)."""
    for name in ("r2_gemm_traffic.json", "r1_gemm_traffic.json"):
        p = os.path.join(ROOT, "profiles", name)
        if os.path.exists(p):
            d = json.load(open(p))
            return d.get("avg_dram_bytes_per_launch"), (f"profiles/{name}: mean of 6 forward launches of one layer pair, "
                                                         f"algorithmic mean {d.get('avg_algorithmic_bytes_per_launch')} B")
    return None, None


def synth_batch(B, device, seed, pin=False):
    """Seeded synthetic hot-path inputs: uni-modal encoder outputs of the shapes RoBERTa-base / CLIP ViT-B/16 give at
    288 px / 32 tokens, ragged text masks (valid length uniform in [8,32]), one-hot VQA targets."""
    g = torch.Generator().manual_seed(seed)
    text = torch.randn(B, LT, HID, generator=g)
    image = torch.randn(B, LI, HID, generator=g)
    lens = torch.randint(8, LT + 1, (B,), generator=g)
    masks = (torch.arange(LT)[None, :] < lens[:, None]).long()
    labels = torch.randint(0, NLAB, (B,), generator=g)
    targets = torch.zeros(B, NLAB)
    targets[torch.arange(B), labels] = 1.0
    out = {"text": text, "image": image, "masks": masks, "targets": targets}
    if pin:
        return {k: v.pin_memory() for k, v in out.items()}
    return {k: v.to(device) for k, v in out.items()}


class ClockSampler(threading.Thread):
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], False

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([t.strip() for t in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        sm = sorted(int(s[0]) for s in self.samples if s[0].isdigit())
        mx = max([int(s[1]) for s in self.samples if s[1].isdigit()], default=None)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for s in self.samples for n, v in zip(names, s[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": reasons, "samples": len(sm)}


# ----------------------------------------------------------------------------- CPU legs: the reference's own code
class _Tok:
    cls_token_id, sep_token_id, eos_token_id, pad_token_id = 101, 102, None, 0


def _ref_available():
    from oracle import ref_shim
    return ref_shim.available()


def reference_fusion_stepper(B, dropout_p=0.1, with_optimizer=True):
    """One fine-tune step of the hot path on the host cores.  Returns (step_fn, kind, description).
    kind "reference": the reference's own M3AETransformerSS (m3ae_module.py:203-312 fusion section, uni-modal encoder
    outputs plugged in as constants) + vqa_head + BCE (objectives.py:201) + backward + AdamW(eps 1e-8, betas (0.9, 0.98)),
    imported from oracle/_ref (or /root/reference) through oracle/ref_shim.py.  kind "port": oracle/m3ae_oracle.py."""
    from oracle import m3ae_oracle as O
    inp = O.make_fusion_inputs(1, B=B)
    if _ref_available():
        from oracle import ref_shim
        mm3, Cls = ref_shim.load_m3ae_module()
        torch.manual_seed(0)
        model = Cls(ref_shim.m3ae_config(drop_rate=dropout_p)).train()
        model.load_state_dict(O.make_fusion_state(0), strict=False)

        class Const(torch.nn.Module):
            def __init__(self, value):
                super().__init__()
                self.value = value

            def forward(self, *a, **k):
                return self.value

        model.language_encoder.embeddings = Const(inp["text_feats"])
        model.vision_encoder = Const(inp["image_feats"])
        batch = {"image": [torch.zeros(B, 3, 288, 288)], "text_ids": torch.zeros(B, 32, dtype=torch.long),
                 "text_labels": torch.full((B, 32), -100), "text_masks": inp["text_masks"]}
        params = [p for p in model.parameters() if p.requires_grad]
        opt = torch.optim.AdamW(params, lr=5e-6, betas=(0.9, 0.98), eps=1e-8, weight_decay=0.01) if with_optimizer else None

        def step():
            for p in params:
                p.grad = None
            ret = model.infer(batch)
            logits = model.vqa_head(ret["multi_modal_cls_feats"])
            loss = torch.nn.functional.binary_cross_entropy_with_logits(logits, inp["targets"]) * inp["targets"].shape[1]
            loss.backward()
            if opt is not None:
                opt.step()

        return step, "reference", "the reference's own M3AETransformerSS fusion section + vqa_head + BCE (oracle/_ref via oracle/ref_shim.py)"
    sd = {k: v.clone().requires_grad_(True) for k, v in O.make_fusion_state(0).items()}
    opt = torch.optim.AdamW(list(sd.values()), lr=5e-6, betas=(0.9, 0.98), eps=1e-8, weight_decay=0.01) if with_optimizer else None

    def step():
        for v in sd.values():
            v.grad = None
        loss, _, _ = O.finetune_step(inp["text_feats"], inp["image_feats"], inp["text_masks"], inp["targets"], sd,
                                     dropout_p=dropout_p, training=True)
        loss.backward()
        if opt is not None:
            opt.step()

    return step, "port", "oracle port (oracle/m3ae_oracle.py; no reference snapshot on this box)"


def cpu_baseline(sample_b=8, steps=3, threads=None):
    """The CPU step of `--impl reference` (same definition, same code) timed inside the `ours` run."""
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    step, kind, what = reference_fusion_stepper(sample_b)
    step()
    best = 1e9
    for _ in range(steps):
        t = time.perf_counter()
        step()
        best = min(best, time.perf_counter() - t)
    return {"value": round(sample_b / best, 3), "unit": "samples/s", "cores": threads, "kind": kind,
            "sample": f"batch {sample_b} (of 64), fwd+bwd+AdamW fp32, dropout 0.1, best of {steps} after 1 warm-up; {what}"}


def cpu_decoder_baseline(B=16, L=16, Lm=359, threads=None):
    """Reference Decoder(max_len=16).search_path (m3ae_decoder.py:141-182: whole-prefix re-run, 6 layers, no cache) at batch B."""
    from oracle import m3ae_oracle as O
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    mem = torch.randn(B, Lm, 768, generator=torch.Generator().manual_seed(3))
    sd = O.make_decoder_state(31)
    if _ref_available():
        from oracle import ref_shim
        ref = ref_shim.load()
        dec = ref.Decoder(6, 768, 8, 3072, 0.1, L, 30522).eval()
        dec.load_state_dict(sd, strict=True)
        kind, what = "reference", "the reference's own Decoder.search_path (oracle/_ref)"

        def run():
            with torch.no_grad():
                return dec.search_path(mem, _Tok)
    else:
        kind, what = "port", "oracle port of Decoder.search_path, all 6 layers evaluated as the reference does"

        def run():
            with torch.no_grad():
                return O.search_path(mem, sd, L, all_layers=True)
    t = time.perf_counter()
    run()
    dt = time.perf_counter() - t
    return {"value": round(B / dt, 3), "unit": "samples/s", "cores": threads, "kind": kind,
            "sample": f"batch {B} (of 128), {L} tokens, memory {Lm} tokens, fp32, one run; {what}"}


def cpu_config1_baseline(B=8, threads=None):
    """BASELINE config 1 as written: the reference's real M3AETransformerSS (its CLIP ViT-B/16 class + random-init HF
    RobertaModel + fusion + vqa_head), infer + vqa_head, eval, CPU fp32, batch 8."""
    if not _ref_available():
        return None
    from oracle import m3ae_oracle as O
    from oracle import ref_shim
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    mm3, Cls = ref_shim.load_m3ae_module(real_encoders=True)
    torch.manual_seed(0)
    model = Cls(ref_shim.m3ae_config()).eval()
    O.seed_model_parameters(model, 5)
    ref_shim.wrap_text_layers(model)
    batch = O.make_model_batch(51, B)

    def run():
        with torch.no_grad():
            return model.vqa_head(model.infer(batch)["multi_modal_cls_feats"])

    run()
    best = 1e9
    for _ in range(2):
        t = time.perf_counter()
        run()
        best = min(best, time.perf_counter() - t)
    return {"value": round(B / best, 3), "unit": "samples/s", "cores": threads, "kind": "reference", "ms": round(best * 1e3, 1),
            "sample": f"batch {B}, infer + vqa_head, eval, fp32, best of 2 after 1 warm-up; the reference's own M3AETransformerSS with its CLIP class + random-init HF RobertaModel"}


# ----------------------------------------------------------------------------- GPU legs
def _event_time(fn, n, sync):
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(i)
    e1.record()
    sync()
    return e0.elapsed_time(e1)


def run_ours(args):
    import torch.distributed as dist

    import m3ae_b200  # noqa: F401
    from m3ae_b200 import ops
    from m3ae_b200.m3ae_module import M3AEFusion

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (impl=ours) needs a B200: the M3AE fusion path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B = args.batch
    dtype = torch.bfloat16 if args.dtype == "bf16" else torch.float32
    torch.manual_seed(0)
    model = M3AEFusion({"drop_rate": 0.1, "precision": 16 if dtype == torch.bfloat16 else 32}, compute_dtype=dtype).to(dev).train()
    params = [p for p in model.parameters()]
    init_probe = params[0].detach().clone()
    opt = torch.optim.AdamW(params, lr=5e-6, betas=(0.9, 0.98), eps=1e-8, weight_decay=0.01, fused=True)
    from m3ae_b200.ddp import GradientAllReducer
    # M3AE_BENCH_FORCE_REDUCER=1: diagnostic — run the N>1 gradient bookkeeping (bucket views, accumulate, zero) on one GPU
    force_reducer = os.environ.get("M3AE_BENCH_FORCE_REDUCER", "0") == "1"
    # fp32 (strict-parity) runs keep the exact fp32 gradient exchange; bf16 runs take the reducer's default (bf16 wire from 8 ranks on)
    reducer = GradientAllReducer(params, world, wire_dtype=torch.float32 if dtype == torch.float32 else None) if (world > 1 or force_reducer) else None
    if reducer is not None:
        reducer.broadcast_parameters(0)

    def step(b):
        if reducer is not None:
            reducer.zero_grad()
        else:
            opt.zero_grad(set_to_none=True)
        loss, _, _ = model.vqa_loss_from_feats(b["text"], b["image"], b["masks"], b["targets"])
        loss.backward()
        if reducer is not None:
            reducer.wait()
        opt.step()
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(n):
            fn(i)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = t.item()
        return ms

    # ---- N > 1: numeric check of the gradient exchange on the GPUs (dropout off): the reducer's averaged gradients of the
    # per-rank batches must equal the gradient of the concatenated batch computed locally without any exchange.
    grad_check = None
    if world > 1:
        model.eval()
        gb = 4
        allb = [synth_batch(gb, dev, 500 + r) for r in range(world)]
        reducer.zero_grad()
        loss, _, _ = model.vqa_loss_from_feats(allb[rank]["text"], allb[rank]["image"], allb[rank]["masks"], allb[rank]["targets"])
        loss.backward()
        reducer.wait()
        torch.cuda.synchronize()
        got = [p.grad.detach().float().clone() for p in params]
        for p in params:
            p.grad = None
        cat = {k: torch.cat([b[k] for b in allb]) for k in allb[0]}
        reducer.enabled = False  # local whole-batch gradients: no exchange
        loss, _, _ = model.vqa_loss_from_feats(cat["text"], cat["image"], cat["masks"], cat["targets"])
        loss.backward()
        torch.cuda.synchronize()
        reducer.enabled = True
        num = torch.sqrt(sum(((g - p.grad.float()) ** 2).sum() for g, p in zip(got, params)))
        den = torch.sqrt(sum((p.grad.float() ** 2).sum() for p in params))
        err = torch.tensor([(num / den).item()], device=dev)
        dist.all_reduce(err, op=dist.ReduceOp.MAX)
        grad_check = {"rel_err_vs_whole_batch": round(err.item(), 6), "per_rank_batch": gb, "gate": 2e-2,
                      "what": "||avg-reduced grads of per-rank batches - local grads of the concatenated batch|| / ||.||, all parameters, max over ranks, dropout off, bf16"}
        if not err.item() < 2e-2:
            raise RuntimeError(f"gradient exchange check failed: {grad_check}")
        for p in params:
            p.grad = None
        reducer.zero_grad()
        model.train()

    # ---- device-resident throughput
    batches = [synth_batch(B, dev, 1000 + rank * 100 + i) for i in range(2)]
    for i in range(max(args.warmup, 3)):
        step(batches[i % 2])
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    l0 = ops.launch_count()
    ms = timed(lambda i: step(batches[i % 2]), args.steps)
    launches = ops.launch_count() - l0
    host_issue = []
    for i in range(3):  # CPU time to ISSUE one step into an idle stream (no sync inside): launch-bound if ~ ms/step
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        step(batches[i % 2])
        host_issue.append((time.perf_counter() - t0) * 1e3)
    torch.cuda.synchronize()
    host_issue_ms = min(host_issue)
    # ---- end-to-end: pinned host inputs -> H2D -> step -> loss D2H, every step.  The H2D of step i+1 is issued on a
    # copy stream while step i runs (double-buffered device inputs); every step's copy is inside the timed region.
    host = [synth_batch(B, dev, 2000 + rank * 100 + i, pin=True) for i in range(2)]
    h2d_bytes = sum(v.numel() * v.element_size() for v in host[0].values())
    loss_host = torch.zeros((), pin_memory=True)
    copy_stream = torch.cuda.Stream()
    dbuf = [{k: torch.empty(v.shape, dtype=v.dtype, device=dev) for k, v in host[0].items()} for _ in range(2)]
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    consumed = [torch.cuda.Event(), torch.cuda.Event()]

    def prefetch(i):
        s = i % 2
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[s])  # the step that last read this slot has finished
            for k, v in host[i % 2].items():
                dbuf[s][k].copy_(v, non_blocking=True)
            ready[s].record(copy_stream)

    def e2e_step(i):
        s = i % 2
        if i == 0:
            prefetch(0)
        prefetch(i + 1)  # next step's inputs travel while this step computes
        torch.cuda.current_stream().wait_event(ready[s])
        loss = step(dbuf[s])
        consumed[s].record()
        loss_host.copy_(loss.detach(), non_blocking=False)  # D2H read of the step's result (synchronises)

    for ev in consumed:
        ev.record()
    e2e_step(0)
    torch.cuda.synchronize()
    for ev in consumed:
        ev.record()
    ms_e2e = timed(e2e_step, args.steps)
    torch.cuda.synchronize()
    if sampler:
        sampler.stop_flag = True
        sampler.join(timeout=2)

    # ---- roofline of the dominant kernel (tcgen05 GEMM): instrumented pass, CUDA events around every GEMM launch
    # EVERY rank runs these steps: each step launches the gradient all-reduce, so a rank-0-only pass deadlocks N > 1
    # (rank 0 inside the collective, the others in the barrier below).  Only rank 0 reports.
    gemm_time_ms, gemm_flops, gemm_n = 0.0, 0.0, 0
    records = []
    import m3ae_b200.functional as Fn
    orig_group, orig_run = ops.gemm_group, Fn._Gemm.run

    def timed_launch(launch, flops):
        a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        launch()
        b_.record()
        records.append((a, b_, flops))

    def group_instrumented(specs, backend=ops.GEMM_AUTO):
        timed_launch(lambda: orig_group(specs, backend), sum(2.0 * s.problem.M * s.problem.N * s.problem.K for s in specs))

    def run_instrumented(self, be, stream):
        timed_launch(lambda: orig_run(self, be, stream), sum(2.0 * q.M * q.N * q.K for q in self.arr))

    ops.gemm_group = group_instrumented  # looked up on the module at call time
    Fn._Gemm.run = run_instrumented       # cached launches of the fused layer node
    try:
        n_prof = min(args.steps, 3)
        for i in range(n_prof):
            step(batches[i % 2])
        torch.cuda.synchronize()
    finally:
        ops.gemm_group, Fn._Gemm.run = orig_group, orig_run
    gemm_time_ms = sum(a.elapsed_time(b_) for a, b_, _ in records)
    gemm_flops = sum(f for _, _, f in records)
    gemm_n = len(records) // max(1, n_prof)
    barrier()
    # N > 1: every rank started from the same seed and saw different batches, so the replicas are still bit-identical
    # if and only if every step applied the same (averaged) gradients — the GPU-side check of the gradient exchange.
    in_sync = None
    if world > 1:
        with torch.no_grad():
            digest = torch.stack([p.detach().double().sum() for p in params] + [p.detach().double().abs().sum() for p in params])
        lo, hi = digest.clone(), digest.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        in_sync = bool(torch.equal(lo, hi))
        moved = bool((params[0].detach() != init_probe).any().item())  # and the optimizer did move the weights
        if not (in_sync and moved):
            raise RuntimeError(f"data-parallel replicas diverged or did not train (in_sync={in_sync}, moved={moved}): the gradient exchange is broken")

    line = None
    if rank == 0:
        pk = peaks()
        sps = world * B * args.steps / (ms / 1e3)
        sps_e2e = world * B * args.steps / (ms_e2e / 1e3)
        achieved = gemm_flops / (gemm_time_ms / 1e3) / 1e12 if gemm_time_ms > 0 else None
        stack_tflops = (sps / world) * FLOPS_FWD_BWD_PER_SAMPLE / 1e12
        line = {
            "metric": "fine-tune samples/sec (M3AE fusion-stack fine-tune step, fwd+bwd+AdamW)",
            "value": round(sps, 2), "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": round(ms / args.steps, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": args.dtype, "data": "synthetic (seeded uni-modal features 288px/32tok, random-init weights)",
            "config": {"workload": "M3AE VQA-RAD fine-tune step (fwd+bwd), fusion stack 6+6 BertCrossLayer + heads + BCE, batch 64/GPU, 288px (325 image tokens) / 32 text tokens",
                       "per_gpu_batch": B, "global_batch": B * world, "dropout": 0.1, "optimizer": "AdamW(fused) inside the timed step",
                       "parallelism": f"dp{world}" + (f" (NCCL gradient all-reduce, {'bf16' if reducer.wire_dtype == torch.bfloat16 else 'fp32'} wire)" if world > 1 else ""),
                       "l2": "no explicit flush: per-step working set (>4 GB of activations + 0.7 GB weights) far exceeds the 126 MB L2"},
            "clocks": sampler.summary() if sampler else None,
            "e2e": {"value": round(sps_e2e, 2), "unit": "samples/s", "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": 4,
                    "ms_per_step": round(ms_e2e / args.steps, 3),
                    "how": "pinned host inputs, H2D of step i+1 on a copy stream during step i (double-buffered), loss read back every step"},
            "gpu_launches": int(launches), "host_issue_ms_per_step": round(host_issue_ms, 3),
            "replicas_in_sync": in_sync, "grad_check": grad_check,
            "roofline": {"bound": "tensor", "achieved": round(achieved, 1) if achieved else None, "peak": pk["tflops"], "unit": "TFLOP/s",
                         "frac": round(achieved / pk["tflops"], 4) if achieved else None, "traffic": gemm_traffic()[0],
                         "traffic_source": gemm_traffic()[1],
                         "kernel": "gemm_tcgen05_kernel", "launches_per_step": gemm_n, "peak_source": pk["source"],
                         "how": "sum of 2MNK over every GEMM launch of a step / sum of CUDA-event durations (instrumented pass, same stream)"},
            "fusion_stack_tflops": round(stack_tflops, 1),
            "fusion_stack_frac_of_peak": round(stack_tflops / pk["tflops"], 4),
        }
    # free the fusion model before the other workloads
    del model, opt, params, batches, host, dbuf, reducer
    torch.cuda.empty_cache()
    if rank == 0 and world == 1 and not args.fusion_only:
        line["decoder"] = bench_decoder(args, dev)
        torch.cuda.empty_cache()
        line["whole_step"] = bench_whole_step(args, dev)
        torch.cuda.empty_cache()
        line["classifier_forward_b8"] = bench_config1(args, dev)
        torch.cuda.empty_cache()
        line["t5_head"] = bench_t5_head(args, dev)
        torch.cuda.empty_cache()
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(sample_b=8, steps=3)
        print(json.dumps(line), file=_STDOUT, flush=True)
    if world > 1:
        dist.destroy_process_group()


def bench_decoder(args, dev):
    """BASELINE config 3: generative decoder head, greedy decode of up to 16 answer tokens, batch 128 (forward only,
    replicas only -> single GPU).  HBM-bound: per step the live layer's weights (18.9 MB bf16), the vocab projection
    (46.9 MB) and the cross K/V cache (B*359*1536*2 B = 141 MB) are streamed once."""
    from m3ae_b200 import ops
    from m3ae_b200.decoder import Decoder
    B, Lm, L = 128, 359, 16
    torch.manual_seed(0)
    dec = Decoder(6, 768, 8, 3072, 0.1, L, 30522, compute_dtype=torch.bfloat16).to(dev).eval()
    g = torch.Generator().manual_seed(3)
    mems = [torch.randn(B, Lm, 768, generator=g).to(dev) for _ in range(2)]
    steps = max(5, min(args.steps, 20))
    for i in range(max(args.warmup, 3)):
        dec.search_path(mems[i % 2], _Tok)
    torch.cuda.synchronize()
    l0, g0 = ops.launch_count(), dec.graph_replayed_launches
    ms = _event_time(lambda i: dec.search_path(mems[i % 2], _Tok), steps, torch.cuda.synchronize) / steps
    launches = (ops.launch_count() - l0) + (dec.graph_replayed_launches - g0)  # eager calls + kernels replayed from the CUDA graph
    # end to end: memory from pinned host fp32 -> H2D -> greedy program -> ids back on the host
    hmem = [torch.randn(B, Lm, 768, generator=g).pin_memory() for _ in range(2)]
    dmem = [torch.empty(B, Lm, 768, device=dev) for _ in range(2)]
    ids_host = torch.empty(B, L, dtype=torch.int64).pin_memory()
    copy_stream = torch.cuda.Stream()
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    consumed = [torch.cuda.Event(), torch.cuda.Event()]

    def prefetch(i):  # the memory of search i travels on a copy stream (while search i-1 decodes)
        s = i % 2
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[s])
            dmem[s].copy_(hmem[s], non_blocking=True)
            ready[s].record(copy_stream)

    def e2e(i):
        s = i % 2
        if i == 0:
            prefetch(0)
        prefetch(i + 1)
        torch.cuda.current_stream().wait_event(ready[s])
        ids = dec.search_path(dmem[s], _Tok)
        consumed[s].record()
        ids_host.copy_(ids, non_blocking=False)  # D2H read of the decoded ids (synchronises)

    for ev in consumed:
        ev.record()
    e2e(0)
    torch.cuda.synchronize()
    for ev in consumed:
        ev.record()
    ms_e2e = _event_time(e2e, steps, torch.cuda.synchronize) / steps
    pk = peaks()
    # algorithmic bytes per greedy step (bf16): live layer weights + vocab projection + cross K/V cache + self K/V cache
    w_bytes = (3 * 768 * 768 + 768 * 768 + 768 * 768 + 768 * 768 + 2 * 768 * 3072 + 30522 * 768) * 2
    kv_bytes = B * Lm * 1536 * 2
    step_bytes = w_bytes + kv_bytes + B * 8 * 1536 * 2  # + average self cache (8 of 16 positions)
    one_off = B * Lm * 768 * 4 + B * Lm * 768 * 2 * 2 + 1536 * 768 * 2  # memory read (fp32) + cast + cross K/V projection
    achieved = (L * step_bytes + one_off) / (ms * 1e-3) / 1e9
    out = {"metric": "greedy-decoded samples/sec (M3AE generative decoder head, <=16 tokens)", "value": round(B / (ms * 1e-3), 1),
           "unit": "samples/s", "n_gpus": 1, "steps": steps, "ms_per_step": round(ms, 3), "dtype": "bf16",
           "config": {"workload": "M3AE generative decoder head (main_decoder_m3ae) greedy decode, 16 tokens, batch 128, memory 359 tokens (KV-cached, live layer only)",
                      "tokens_per_s": round(B * L / (ms * 1e-3), 1), "us_per_token_step": round(ms * 1e3 / L, 1)},
           "gpu_launches": int(launches),
           "e2e": {"value": round(B / (ms_e2e * 1e-3), 1), "unit": "samples/s", "h2d_bytes_per_step": B * Lm * 768 * 4,
                   "d2h_bytes_per_step": B * L * 8, "ms_per_step": round(ms_e2e, 3),
                   "how": "fp32 memory [128,359,768] from pinned host memory every search (H2D of search i+1 on a copy stream while search i decodes, "
                          "double-buffered), decoded ids read back to the host every search"},
           "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": pk["hbm_gbs"], "unit": "GB/s",
                        "frac": round(achieved / pk["hbm_gbs"], 4), "traffic": None,
                        "how": "algorithmic bytes per greedy step (weights 65.8 MB + cross K/V 141 MB + self K/V) x 16 + one-off memory projection, / CUDA-event time of search_path"}}
    if not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_decoder_baseline()
    del dec, mems
    return out


def bench_whole_step(args, dev):
    """BASELINE config 2 with the uni-modal encoders inside the step: M3AETransformerSS.training_step (CLIP ViT-B/16 +
    RoBERTa-base + fusion + vqa_head + BCE) -> backward -> AdamW over all trainable parameters, batch 64, bf16, dropout 0.1.
    Algorithmic work 309.7 GFLOP per sample fwd+bwd."""
    from m3ae_b200.m3ae_module import M3AETransformerSS
    from oracle import m3ae_oracle as O  # synthetic batch-dict generator only
    B = args.batch
    torch.manual_seed(0)
    model = M3AETransformerSS({"drop_rate": 0.1}, compute_dtype=torch.bfloat16).to(dev).train()
    opt = torch.optim.AdamW([p for p in model.parameters()], lr=1e-5, betas=(0.9, 0.98), eps=1e-8, weight_decay=0.01, fused=True)
    batches = [O.make_model_batch(70 + i, B, device=dev) for i in range(2)]

    def step(i):
        opt.zero_grad(set_to_none=True)
        loss = model.training_step(batches[i % 2], i)
        loss.backward()
        opt.step()

    for i in range(3):
        step(i)
    steps = max(3, min(args.steps, 10))
    ms = _event_time(step, steps, torch.cuda.synchronize) / steps
    pk = peaks()
    tf = B / (ms * 1e-3) * 309.7e9 / 1e12
    out = {"metric": "fine-tune samples/sec (whole M3AE: CLIP ViT-B/16 + RoBERTa-base + fusion + vqa_head, fwd+bwd+AdamW)",
           "value": round(B / (ms * 1e-3), 1), "unit": "samples/s", "ms_per_step": round(ms, 3), "steps": steps, "dtype": "bf16",
           "config": {"workload": "M3AETransformerSS.training_step on a synthetic batch dict (images 288x288, 32 tokens), batch 64, dropout 0.1, AdamW(fused) over all 322 M parameters"},
           "model_tflops": round(tf, 1), "frac_of_peak": round(tf / pk["tflops"], 4),
           "how": "CUDA events around K steps; 309.7 GFLOP per sample fwd+bwd (ViT 54.56 + RoBERTa 5.47 + fusion 42.78 + heads 0.43, x3)"}
    del model, opt, batches
    return out


def bench_config1(args, dev):
    """BASELINE config 1: classifier forward (infer + vqa_head), batch 8, 288 px / 32 tokens, eval.  GPU bf16 through the
    drop-in next to the reference's CPU fp32 time (parity of this exact configuration: tests/test_full_model_gpu.py)."""
    from m3ae_b200.m3ae_module import M3AETransformerSS
    from oracle import m3ae_oracle as O
    torch.manual_seed(0)
    model = M3AETransformerSS({"drop_rate": 0.0}, compute_dtype=torch.bfloat16)
    O.seed_model_parameters(model, 5)
    model = model.to(dev).eval()
    batch = O.make_model_batch(51, 8, device=dev)

    def run(i):
        with torch.no_grad():
            return model.vqa_head(model.infer(batch)["multi_modal_cls_feats"])

    for i in range(3):
        run(i)
    ms = _event_time(run, 20, torch.cuda.synchronize) / 20
    out = {"metric": "classifier forward samples/sec (infer + vqa_head, batch 8)", "value": round(8 / (ms * 1e-3), 1), "unit": "samples/s",
           "ms_per_forward": round(ms, 3), "dtype": "bf16", "model_tflops": round(8 / (ms * 1e-3) * 103.24e9 / 1e12, 1)}

    def run_graphed(i):  # the same forward replayed from a CUDA graph (M3AETransformerSS.vqa_logits_graphed): input copies + replay + logits clone
        with torch.no_grad():
            return model.vqa_logits_graphed(batch)

    for i in range(3):
        run_graphed(i)
    ms_g = _event_time(run_graphed, 20, torch.cuda.synchronize) / 20
    out["cuda_graph"] = {"value": round(8 / (ms_g * 1e-3), 1), "unit": "samples/s", "ms_per_forward": round(ms_g, 3),
                         "kernels_per_replay": next(iter(model._infer_graphs.values()))["kernels"]}
    del model
    if not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_config1_baseline()
    return out


T5_ANSWERS = ["no", "yes", "left lower lobe pneumonia", "right upper lobe", "abdomen", "chest x ray", "the liver", "pleural effusion"]


def _t5_config():
    """finetune_m3ae.sh:9-16: t5_max_length 12, cls feature only, top 4 encoder / decoder blocks unfrozen."""
    return {"drop_rate": 0.1, "hidden_size": 768, "load_path_t5": "", "t5_max_length": 12, "mm_encoder_inputs_include_cls_feats": True,
            "mm_encoder_inputs_include_imagetext_feats": False, "mm_encoder_inputs_mm_feats_width": 0,
            "loss_names": {"mlm": 0, "mim": 0, "itm": 0, "vqa": 1, "cls": 0, "irtr": 0}}


def _t5_batch(O, seed, B, device):
    batch = O.make_model_batch(seed, B, device=device)
    batch["text"] = ["what abnormality is seen in this image"] * B
    batch["vqa_answer"] = [[T5_ANSWERS[i % len(T5_ANSWERS)]] for i in range(B)]
    return batch


def cpu_t5_baseline(B=4, threads=None):
    """The reference's own T5VQA_MMEncoderInput (m3ae_t5_mm_encoder_input.py:191-295, oracle/_ref) with a random-init t5-small,
    one teacher-forced fine-tune step on the host cores: forward (frozen M3AE + T5) + backward + AdamW.  Its per-step beam-4
    generation for metrics (:250-259) is NOT part of "teacher-forced" and is switched off through max_length 1."""
    if not _ref_available():
        return None
    from oracle import m3ae_oracle as O
    from oracle import ref_shim
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    Cls = ref_shim.load_t5_head(O.SynthT5Tokenizer())
    cfg = ref_shim.m3ae_config(drop_rate=0.1)
    cfg.update(_t5_config())
    torch.manual_seed(0)
    model = Cls(cfg).train()
    ref_shim.wrap_text_layers(model.m3ae)
    model.unfreeze_top_layers(4, 4)
    model.current_tasks = ["vqa"]
    model.max_answer_length = 2  # shortest generation the reference's forward allows: the metrics-only beam search is not timed work
    params = [p for p in model.parameters() if p.requires_grad]
    opt = torch.optim.AdamW(params, lr=1e-5, betas=(0.9, 0.98), eps=1e-8, weight_decay=0.01)
    batch = _t5_batch(O, 90, B, "cpu")

    def step():
        opt.zero_grad(set_to_none=True)
        model(batch)["vqa_loss"].backward()
        opt.step()

    step()
    t = time.perf_counter()
    step()
    dt = time.perf_counter() - t
    return {"value": round(B / dt, 3), "unit": "samples/s", "cores": threads, "kind": "reference", "ms": round(dt * 1e3, 1),
            "sample": f"batch {B} (of 64), one step after 1 warm-up, fp32, dropout 0.1; the reference's own T5VQA_MMEncoderInput.forward (train branch, "
                      "generation for metrics cut to 2 tokens) + backward + AdamW over the unfrozen top-4 T5 blocks, random-init t5-small"}


def bench_t5_head(args, dev):
    """BASELINE config 4: T5 generative head, teacher-forced fine-tune step, bf16, batch 64: frozen M3AE.infer on the sm_100a kernels
    (dropout active: the frozen M3AE stays in train mode, as in the reference) -> prepare_inputs -> HF T5-small encoder (512 positions)
    -> teacher-forced decoder loss -> backward through the unfrozen top-4 encoder / decoder blocks -> AdamW.  Reported as samples/s
    only: the T5 arithmetic is third-party library code on both sides."""
    from m3ae_b200.t5_head import T5VQA_MMEncoderInput
    from oracle import m3ae_oracle as O  # synthetic batch-dict generator only
    B = args.batch
    torch.manual_seed(0)
    model = T5VQA_MMEncoderInput(_t5_config(), compute_dtype=torch.bfloat16, generate_for_metrics=False).to(dev).train()
    model.unfreeze_top_layers(4, 4)
    model.current_tasks = ["vqa"]
    params = [p for p in model.parameters() if p.requires_grad]
    opt = torch.optim.AdamW(params, lr=1e-5, betas=(0.9, 0.98), eps=1e-8, weight_decay=0.01, fused=True)
    batches = [_t5_batch(O, 90 + i, B, dev) for i in range(2)]

    def step(i):
        opt.zero_grad(set_to_none=True)
        model(batches[i % 2])["vqa_loss"].backward()
        opt.step()

    for i in range(3):
        step(i)
    steps = max(3, min(args.steps, 10))
    ms = _event_time(step, steps, torch.cuda.synchronize) / steps
    out = {"metric": "fine-tune samples/sec (M3AE T5 generative head, teacher-forced)", "value": round(B / (ms * 1e-3), 1), "unit": "samples/s",
           "ms_per_step": round(ms, 3), "steps": steps, "dtype": "bf16",
           "config": {"workload": "T5VQA_MMEncoderInput teacher-forced fine-tune step (main_t5_m3ae), batch 64, 288 px / 32 tokens, cls feature as encoder input "
                                  "(512 positions), top-4 T5 blocks trainable, AdamW(fused); frozen M3AE forward on the sm_100a kernels, T5-small = HF library code under bf16 autocast",
                      "trainable_params": int(sum(p.numel() for p in params))}}
    del model, opt, batches
    if not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_t5_baseline()
    return out


def run_decoder(args):
    import m3ae_b200  # noqa: F401
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    out = bench_decoder(args, dev)
    out.update({"warmup": max(args.warmup, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "data": "synthetic (seeded memory [128,359,768], random-init weights)"})
    print(json.dumps(out), file=_STDOUT, flush=True)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    B = 8
    steps = max(1, min(args.steps, 5))
    threads = os.cpu_count()
    torch.set_num_threads(threads)
    one, kind, what = reference_fusion_stepper(B)
    warm = max(1, min(args.warmup, 2))
    for _ in range(warm):
        one()
    t = time.perf_counter()
    for _ in range(steps):
        one()
    dt = time.perf_counter() - t
    sps = B * steps / dt
    sample = f"bounded sample: batch {B} (of 64) x {steps} steps, fp32, dropout 0.1, AdamW step included, {threads} host threads; {what}"
    print(json.dumps({
        "impl": "reference", "metric": "fine-tune samples/sec (M3AE fusion-stack fine-tune step, fwd+bwd+AdamW)",
        "value": round(sps, 3), "unit": "samples/s", "n_gpus": args.gpus, "steps": steps, "warmup": warm,
        "ms_per_step": round(dt / steps * 1e3, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic (seeded uni-modal features 288px/32tok, random-init weights)",
        "config": {"workload": "M3AE VQA-RAD fine-tune step (fwd+bwd), fusion stack 6+6 BertCrossLayer + heads + BCE, 288px / 32 tokens", "per_gpu_batch": B},
        "cpu_baseline": {"value": round(sps, 3), "unit": "samples/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": round(sps, 3), "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), file=_STDOUT, flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=64, help="per-GPU batch (BASELINE config[1]: 64)")
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "f32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--fusion-only", action="store_true", help="skip the decoder / whole-step / config-1 objects of the N=1 line")
    ap.add_argument("--watchdog", type=int, default=1500,
                    help="seconds after which a stuck rank dumps every thread's Python stack to stderr and exits 1 (0 = off)")
    ap.add_argument("--workload", default="finetune", choices=["finetune", "decoder"],
                    help="finetune = BASELINE config[1] (default, the headline); decoder = config[2] greedy decoding only")
    args = ap.parse_args()
    if args.watchdog > 0:  # a hung collective or kernel must fail loudly, not hang the caller
        import faulthandler
        faulthandler.dump_traceback_later(args.watchdog, exit=True)
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "decoder":
        run_decoder(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
