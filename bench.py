#!/usr/bin/env python
"""Benchmark of the M3AE fusion fine-tune hot path on B200 (BASELINE.json config[1]).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one fine-tune pass of the hot path over one synthetic batch of 64 samples per GPU:
modality projections -> 6+6 co-attention layers (fwd) -> poolers -> vqa_head -> BCE loss -> backward through all of
it -> (N>1: NCCL gradient all-reduce) -> AdamW step.  Dropout (p=0.1) is active as in the reference's fine-tune.
`value` is samples/s with inputs resident in HBM; `e2e` re-measures the same step driven from pinned HOST buffers
(H2D of the step's inputs + D2H of the loss inside the timed region).  `roofline` is the tcgen05 GEMM kernel
(dominant kernel): algorithmic FLOPs of every GEMM launch in a step / summed CUDA-event durations, against the
measured bf16 peak in MEASURED_PEAKS.json.  `cpu_baseline` / `--impl reference` time the oracle port of the
reference's CPU path (the reference is pure Python+torch and cannot travel to the GPU box; oracle/m3ae_oracle.py is
pinned to it by golden vectors) on the host cores.
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
    p = os.path.join(ROOT, "profiles", "r1_gemm_traffic.json")
    if not os.path.exists(p):
        return None, None
    d = json.load(open(p))
    return d.get("avg_dram_bytes_per_launch"), (f"profiles/r1_gemm_traffic.json: mean of 6 forward launches of one layer pair, "
                                                 f"algorithmic mean {d.get('avg_algorithmic_bytes_per_launch')} B")


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
    reducer = GradientAllReducer(params, world) if (world > 1 or force_reducer) else None

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
    # ---- end-to-end: pinned host inputs -> H2D -> step -> loss D2H, every step
    host = [synth_batch(B, dev, 2000 + rank * 100 + i, pin=True) for i in range(2)]
    h2d_bytes = sum(v.numel() * v.element_size() for v in host[0].values())
    loss_host = torch.zeros((), pin_memory=True)

    def e2e_step(i):
        b = {k: v.to(dev, non_blocking=True) for k, v in host[i % 2].items()}
        loss = step(b)
        loss_host.copy_(loss.detach(), non_blocking=False)  # D2H read of the step's result (synchronises)

    e2e_step(0)
    ms_e2e = timed(e2e_step, args.steps)
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
                       "parallelism": f"dp{world}" + (" (NCCL gradient all-reduce)" if world > 1 else ""),
                       "l2": "no explicit flush: per-step working set (>4 GB of activations + 0.7 GB weights) far exceeds the 126 MB L2"},
            "clocks": sampler.summary() if sampler else None,
            "e2e": {"value": round(sps_e2e, 2), "unit": "samples/s", "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": 4,
                    "ms_per_step": round(ms_e2e / args.steps, 3)},
            "gpu_launches": int(launches), "host_issue_ms_per_step": round(host_issue_ms, 3),
            "replicas_in_sync": in_sync,
            "roofline": {"bound": "tensor", "achieved": round(achieved, 1) if achieved else None, "peak": pk["tflops"], "unit": "TFLOP/s",
                         "frac": round(achieved / pk["tflops"], 4) if achieved else None, "traffic": gemm_traffic()[0],
                         "traffic_source": gemm_traffic()[1],
                         "kernel": "gemm_tcgen05_kernel", "launches_per_step": gemm_n, "peak_source": pk["source"],
                         "how": "sum of 2MNK over every GEMM launch of a step / sum of CUDA-event durations (instrumented pass, same stream)"},
            "fusion_stack_tflops": round(stack_tflops, 1),
            "fusion_stack_frac_of_peak": round(stack_tflops / pk["tflops"], 4),
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(sample_b=8, steps=3)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_decoder(args):
    """BASELINE config 3: generative decoder head, greedy decode of up to 16 answer tokens, batch 128 (forward only,
    replicas only -> single GPU).  HBM-bound: per step the live layer's weights (18.9 MB bf16), the vocab projection
    (46.9 MB) and the cross K/V cache (B*359*1536*2 B = 141 MB) are streamed once."""
    import m3ae_b200  # noqa: F401
    from m3ae_b200 import ops
    from m3ae_b200.decoder import Decoder
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    B, Lm, L = 128, 359, 16
    torch.manual_seed(0)
    dec = Decoder(6, 768, 8, 3072, 0.1, L, 30522, compute_dtype=torch.bfloat16).to(dev).eval()

    class Tok:
        cls_token_id, sep_token_id, eos_token_id, pad_token_id = 101, 102, None, 0

    g = torch.Generator().manual_seed(3)
    mems = [torch.randn(B, Lm, 768, generator=g).to(dev) for _ in range(2)]
    for i in range(max(args.warmup, 3)):
        dec.search_path(mems[i % 2], Tok)
    torch.cuda.synchronize()
    l0, g0 = ops.launch_count(), dec.graph_replayed_launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        ids = dec.search_path(mems[i % 2], Tok)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    launches = (ops.launch_count() - l0) + (dec.graph_replayed_launches - g0)  # eager calls + kernels replayed from the CUDA graph
    pk = peaks()
    # algorithmic bytes per greedy step (bf16): live layer weights + vocab projection + cross K/V cache + self K/V cache
    w_bytes = (3 * 768 * 768 + 768 * 768 + 768 * 768 + 768 * 768 + 2 * 768 * 3072 + 30522 * 768) * 2
    kv_bytes = B * Lm * 1536 * 2
    step_bytes = w_bytes + kv_bytes + B * 8 * 1536 * 2  # + average self cache (8 of 16 positions)
    one_off = B * Lm * 768 * 4 + B * Lm * 768 * 2 * 2 + 1536 * 768 * 2  # memory read (fp32) + cast + cross K/V projection
    achieved = (L * step_bytes + one_off) / (ms * 1e-3) / 1e9
    line = {"metric": "greedy-decoded samples/sec (M3AE generative decoder head, <=16 tokens)", "value": round(B / (ms * 1e-3), 1),
            "unit": "samples/s", "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": round(ms, 3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic (seeded memory [128,359,768], random-init weights)",
            "config": {"workload": "M3AE generative decoder head greedy decode, 16 tokens, batch 128, memory 359 tokens (KV-cached, live layer only)",
                       "tokens_per_s": round(B * L / (ms * 1e-3), 1), "us_per_token_step": round(ms * 1e3 / L, 1)},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": pk["hbm_gbs"], "unit": "GB/s",
                         "frac": round(achieved / pk["hbm_gbs"], 4), "traffic": None,
                         "how": "algorithmic bytes per greedy step (weights 65.8 MB + cross K/V 141 MB + self K/V) x 16 + one-off memory projection, / CUDA-event time of search_path"}}
    print(json.dumps(line), flush=True)


def cpu_baseline(sample_b=8, steps=3, threads=None):
    """Oracle port of the reference CPU path (fusion fwd + vqa head + BCE + backward through everything) on the host."""
    from oracle import m3ae_oracle as O
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    sd = {k: v.clone().requires_grad_(True) for k, v in O.make_fusion_state(0).items()}
    inp = O.make_fusion_inputs(1, B=sample_b)

    def one():
        for v in sd.values():
            v.grad = None
        loss, _, _ = O.finetune_step(inp["text_feats"], inp["image_feats"], inp["text_masks"], inp["targets"], sd,
                                     dropout_p=0.1, training=True)
        loss.backward()

    one()
    best = 1e9
    for _ in range(steps):
        t = time.perf_counter()
        one()
        best = min(best, time.perf_counter() - t)
    return {"value": round(sample_b / best, 3), "unit": "samples/s", "cores": threads, "kind": "port",
            "sample": f"batch {sample_b} (of 64), fwd+bwd fp32 with dropout 0.1, no optimizer step, best of {steps} after 1 warm-up; torch CPU ops via oracle/m3ae_oracle.py"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    B = 8
    steps = max(1, min(args.steps, 5))
    from oracle import m3ae_oracle as O
    threads = os.cpu_count()
    torch.set_num_threads(threads)
    sd = {k: v.clone().requires_grad_(True) for k, v in O.make_fusion_state(0).items()}
    opt = torch.optim.AdamW(list(sd.values()), lr=5e-6, betas=(0.9, 0.98), eps=1e-8, weight_decay=0.01)
    inp = O.make_fusion_inputs(1, B=B)

    def one():
        opt.zero_grad(set_to_none=True)
        loss, _, _ = O.finetune_step(inp["text_feats"], inp["image_feats"], inp["text_masks"], inp["targets"], sd,
                                     dropout_p=0.1, training=True)
        loss.backward()
        opt.step()

    for _ in range(max(1, min(args.warmup, 2))):
        one()
    t = time.perf_counter()
    for _ in range(steps):
        one()
    dt = time.perf_counter() - t
    sps = B * steps / dt
    sample = f"bounded sample: batch {B} (of 64) x {steps} steps, fp32, dropout 0.1, AdamW step included, {threads} host threads; oracle port (reference is pure Python/torch and absent on this box)"
    print(json.dumps({
        "impl": "reference", "metric": "fine-tune samples/sec (M3AE fusion-stack fine-tune step, fwd+bwd+AdamW)",
        "value": round(sps, 3), "unit": "samples/s", "n_gpus": args.gpus, "steps": steps, "warmup": max(1, min(args.warmup, 2)),
        "ms_per_step": round(dt / steps * 1e3, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic (seeded uni-modal features 288px/32tok, random-init weights)",
        "config": {"workload": "M3AE VQA-RAD fine-tune step (fwd+bwd), fusion stack 6+6 BertCrossLayer + heads + BCE, 288px / 32 tokens", "per_gpu_batch": B},
        "cpu_baseline": {"value": round(sps, 3), "unit": "samples/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": round(sps, 3), "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=64, help="per-GPU batch (BASELINE config[1]: 64)")
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "f32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--watchdog", type=int, default=1500,
                    help="seconds after which a stuck rank dumps every thread's Python stack to stderr and exits 1 (0 = off)")
    ap.add_argument("--workload", default="finetune", choices=["finetune", "decoder"],
                    help="finetune = BASELINE config[1] (default, the headline); decoder = config[2] greedy decoding")
    args = ap.parse_args()
    if args.watchdog > 0:  # a hung collective or kernel must fail loudly, not hang the caller
        import faulthandler
        faulthandler.dump_traceback_later(args.watchdog, exit=True)
    if args.workload == "decoder":
        run_decoder(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
