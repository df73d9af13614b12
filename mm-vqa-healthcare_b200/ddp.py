"""Data-parallel gradient exchange: the ONLY collective on the M3AE fine-tune path (reference: pytorch_lightning
`distributed_backend="ddp"`, main.py:63, i.e. bucketed NCCL all-reduce of gradients).

One process per GPU.  Parameters' `.grad` tensors are views into a few large flat fp32 buckets (reverse registration
order ~ reverse execution order, so the last layers' bucket completes first); a post-accumulate-grad hook per
parameter launches the bucket's all-reduce (sum of grads pre-scaled by 1/world) on a side stream the moment its last
gradient lands, so the exchange over NVLink/NVSwitch overlaps the remaining backward GEMMs.  No TP/PP/SP: samples
are independent through the whole path.
"""
import torch
import torch.distributed as dist


class GradientAllReducer:
    """Gradient exchange for one-process-per-GPU data parallelism.

    Zero-copy groups.  The fused BertCrossLayer node returns the 26 parameter gradients of a layer as views of ONE flat
    fp32 buffer (functional.py::_backward_streams).  When `.grad` is None, autograd adopts such a view without copying,
    so the reducer all-reduces the layer's buffer IN PLACE as soon as the layer's last gradient has landed: no bucket
    copy, no per-parameter accumulate kernel, no bucket zero fill (those cost 1.4 ms of a 15.2 ms step when every
    gradient was accumulated into pre-assigned bucket views — measured with M3AE_BENCH_FORCE_REDUCER=1).
    Groups are discovered on the first exchange: parameters whose gradients share one storage that they cover exactly.
    Bucketed parameters.  Everything else (heads, poolers, projections; every parameter of a model whose gradients do
    not share storage) keeps `.grad` as a view into flat fp32 buckets that are all-reduced when complete.

    bf16 wire format (CUDA; default from 8 ranks on, M3AE_DDP_WIRE=fp32|bf16|auto or wire_dtype= override).  The payload of an exchange unit is
    cast to bf16 on the communication stream, all-reduced (NCCL AVG) and cast back into the fp32 gradient buffer: half the
    bytes over NVLink; master gradients, optimizer and its state stay fp32 (the trade PyTorch DDP's bf16_compress_hook
    makes).  Measured at N = 2 (profiles/r2_ddp_wire_ab_n2.txt): 15.81 ms/step against 15.71 ms with the exact fp32
    exchange (neutral: the two casts cost what the shorter collective saves; N = 4: 15.60 vs 15.59); at N = 8 (profiles/r2_ddp_wire_ab.txt):
    15.61 against 16.12 ms/step — the collective's kernels occupy SMs and HBM beside the persistent GEMM for half as long
    (in-step GEMM rate 985 vs 958 TFLOP/s).  Averaged gradients differ from the exact exchange by 3.4e-3 relative
    (bench grad_check, gate 2e-2); replicas stay bit-identical.

    Use: `zero_grad()` before every backward, `wait()` after it, then the optimizer step.  One backward per exchange.
    """

    def __init__(self, params, world_size=None, bucket_mb=None, group=None, overlap=None, zero_copy=None, wire_dtype=None):
        """overlap=True: an exchange starts from the grad hook of its last gradient, concurrently with the rest of
        backward; False: everything is exchanged from wait().  Defaults: environment M3AE_DDP_OVERLAP (0/1, default 1),
        M3AE_DDP_BUCKET_MB (default 64; size does not matter between 16 and 128, profiles/r1_ddp_ab_n2.txt),
        M3AE_DDP_ZERO_COPY (0/1, default 1)."""
        import os
        if overlap is None:
            overlap = os.environ.get("M3AE_DDP_OVERLAP", "1") != "0"
        if zero_copy is None:
            zero_copy = os.environ.get("M3AE_DDP_ZERO_COPY", "1") != "0"
        if bucket_mb is None:
            bucket_mb = float(os.environ.get("M3AE_DDP_BUCKET_MB", "64"))
        self.overlap, self.zero_copy, self.bucket_mb = bool(overlap), bool(zero_copy), bucket_mb
        self.params = [p for p in params if p.requires_grad]
        self.world = world_size if world_size is not None else (dist.get_world_size(group) if dist.is_initialized() else 1)
        if wire_dtype is None:
            # auto: exact fp32 exchange below 8 ranks (bf16 measured neutral at N = 2 and N = 4), bf16 wire from 8 ranks on, where
            # the collective's kernels run long enough beside the persistent GEMM to matter (N = 8: 16.12 -> 15.61 ms/step,
            # profiles/r2_ddp_wire_ab.txt).  M3AE_DDP_WIRE=fp32 / bf16 or wire_dtype= force either.
            env = os.environ.get("M3AE_DDP_WIRE", "auto")
            wire_dtype = torch.bfloat16 if env == "bf16" or (env == "auto" and self.world >= 8) else torch.float32
        if not self.params[0].is_cuda:
            wire_dtype = torch.float32  # gloo (CPU tests): exact exchange
        self.wire_dtype = wire_dtype
        self._wire = {}          # exchange unit -> persistent bf16 staging buffer
        self.group = group
        self.cuda = self.params[0].is_cuda
        self.comm_stream = torch.cuda.Stream() if self.cuda else None
        # M3AE_DDP_RESERVE_SMS: SMs the persistent GEMM leaves to NCCL's kernels while an exchange overlaps backward
        # (default 0; A/B in profiles/r2_ddp_reserve_ab_n2.txt)
        reserve = int(os.environ.get("M3AE_DDP_RESERVE_SMS", "0"))
        if self.cuda and self.world > 1 and reserve > 0:
            from . import ops
            ops.set_gemm_reserved_sms(reserve)
        self.buckets = []        # (flat, [params]) — bucketed parameters
        self.groups = []         # [params] — zero-copy groups
        self._slot = {}          # id(param) -> ("b", bucket index) | ("g", group index)
        self._pending, self._launched, self._works, self._writeback = {}, {}, [], []
        self._learned = not self.zero_copy
        self.enabled = True      # False: backward passes do not exchange anything (local reference gradients)
        for p in self.params:
            p.register_post_accumulate_grad_hook(self._hook)
        if self._learned:
            self._build_buckets(self.params)
        self.zero_grad()

    # ------------------------------------------------------------------ construction
    def _build_buckets(self, ps_all):
        cap = int(self.bucket_mb * (1 << 20)) // 4
        cur, cur_n = [], 0
        for p in reversed(ps_all):  # reverse registration order ~ reverse execution order
            if cur and cur_n + p.numel() > cap:
                self._make_bucket(cur)
                cur, cur_n = [], 0
            cur.append(p)
            cur_n += p.numel()
        if cur:
            self._make_bucket(cur)

    def _make_bucket(self, ps):
        n = sum(p.numel() for p in ps)
        flat = torch.zeros(n, dtype=torch.float32, device=ps[0].device)
        off = 0
        for p in ps:
            view = flat[off:off + p.numel()].view_as(p)
            if p.grad is not None:
                view.copy_(p.grad)
            p.grad = view
            self._slot[id(p)] = ("b", len(self.buckets))
            off += p.numel()
        self.buckets.append((flat, list(ps)))

    def _learn(self):
        """First exchange: find the parameters whose gradients tile one storage exactly (a fused layer's flat buffer)."""
        by_storage = {}
        for p in self.params:
            if p.grad is not None and p.grad.is_contiguous() and p.grad.dtype == torch.float32:
                by_storage.setdefault(p.grad.untyped_storage().data_ptr(), []).append(p)
        grouped = set()
        for ps in by_storage.values():
            if len(ps) >= 2 and sum(p.numel() for p in ps) * 4 == ps[0].grad.untyped_storage().nbytes():
                for p in ps:
                    self._slot[id(p)] = ("g", len(self.groups))
                    grouped.add(id(p))
                self.groups.append(list(ps))
        self._build_buckets([p for p in self.params if id(p) not in grouped])
        self._learned = True

    # ------------------------------------------------------------------ per step
    def zero_grad(self):
        self._works, self._writeback = [], []
        if not self._learned:
            for p in self.params:
                p.grad = None
            return
        for gi, ps in enumerate(self.groups):
            for p in ps:
                p.grad = None  # autograd adopts the layer's flat-buffer views: no copy
            self._pending["g", gi], self._launched["g", gi] = len(ps), False
        for bi, (flat, ps) in enumerate(self.buckets):
            flat.zero_()
            self._pending["b", bi], self._launched["b", bi] = len(ps), False
            off = 0
            for p in ps:  # re-attach views if an optimizer/zero_grad(set_to_none=True) dropped them
                if p.grad is None or p.grad.data_ptr() != flat.data_ptr() + 4 * off:
                    p.grad = flat[off:off + p.numel()].view_as(p)
                off += p.numel()

    def _hook(self, param):
        if not self._learned or not self.enabled:
            return
        key = self._slot[id(param)]
        self._pending[key] -= 1
        if self._pending[key] == 0 and self.overlap:
            self._launched[key] = True
            self._launch(key)

    def _flat_of(self, key):
        """The tensor to all-reduce for an exchange unit, plus a write-back list.  The collective issued for a unit has the
        SAME size on every rank whatever happened locally (a layer without a gradient on this rank, a gradient layout
        that differs from the first step): ranks whose collective sequences diverge hang NCCL."""
        kind, i = key
        if kind == "b":
            return self.buckets[i][0], None
        group = self.groups[i]
        ps = [p for p in group if p.grad is not None]
        if len(ps) == len(group):
            st = ps[0].grad.untyped_storage()
            same = all(p.grad.untyped_storage().data_ptr() == st.data_ptr() for p in ps)
            if same and sum(p.numel() for p in ps) * 4 == st.nbytes():
                return torch.empty(0, dtype=torch.float32, device=ps[0].device).set_(st), None  # the layer's buffer, in place
        # no gradient here, or a different layout: stage through a flat buffer of the group's size (zeros where absent)
        flat = torch.zeros(sum(p.numel() for p in group), dtype=torch.float32, device=group[0].device)
        back, off = [], 0
        for p in group:
            view = flat[off:off + p.numel()].view_as(p)
            if p.grad is not None:
                view.copy_(p.grad)
            back.append((p, view))
            off += p.numel()
        return flat, back

    def _launch(self, key):
        if self.world == 1:
            return
        flat, back = self._flat_of(key)
        if self.cuda:
            self.comm_stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.comm_stream):  # NCCL averages inside the reduction: no separate pre-scale pass
                if self.wire_dtype == torch.float32:
                    self._works.append(dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=self.group, async_op=True))
                else:
                    from . import ops  # the library's vectorised cast kernels, issued on the communication stream
                    wire = self._wire.get(key)
                    if wire is None or wire.numel() != flat.numel():
                        wire = self._wire[key] = torch.empty(flat.numel(), dtype=self.wire_dtype, device=flat.device)
                    ops.cast(flat, wire)
                    dist.all_reduce(wire, op=dist.ReduceOp.AVG, group=self.group, async_op=True).wait()  # stream-ordered on comm_stream
                    ops.cast(wire, flat)
        else:
            flat.mul_(1.0 / self.world)
            self._works.append(dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True))
        if back is not None:
            self._writeback.append(back)

    def wait(self):
        """Call after backward(): starts every exchange that has not started (first step, overlap=False, parameters
        without a gradient) and makes the compute stream wait for all of them."""
        if not self._learned:
            self._learn()
            for gi in range(len(self.groups)):
                self._launched["g", gi] = False
            for bi in range(len(self.buckets)):
                self._launched["b", bi] = False
        for key in [("g", i) for i in range(len(self.groups))] + [("b", i) for i in range(len(self.buckets))]:
            if not self._launched.get(key, False):
                self._launched[key] = True
                self._launch(key)
        for w in self._works:
            w.wait()
        if self.cuda and self.world > 1:
            torch.cuda.current_stream().wait_stream(self.comm_stream)
        for back in self._writeback:  # staged groups: hand the averaged values to the parameters
            for p, view in back:
                if p.grad is None:
                    p.grad = view
                else:
                    p.grad.copy_(view)
        self._works, self._writeback = [], []

    def broadcast_parameters(self, src=0):
        """What PL's DDP does at start-up: every replica takes rank `src`'s parameters (and the caller's buffers stay as
        they are).  Replicas that were built from the same seed need it only as a safety net."""
        if self.world > 1:
            with torch.no_grad():
                for p in self.params:
                    dist.broadcast(p.data, src=src, group=self.group)

    def payload_bytes(self):
        return sum(p.numel() for p in self.params) * (4 if self.wire_dtype == torch.float32 else 2)
