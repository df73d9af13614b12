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
    def __init__(self, params, world_size=None, bucket_mb=64, group=None):
        self.params = [p for p in params if p.requires_grad]
        self.world = world_size if world_size is not None else (dist.get_world_size(group) if dist.is_initialized() else 1)
        self.group = group
        self.buckets = []  # (flat, [params])
        self._pending = {}
        self._works = []
        cap = bucket_mb * (1 << 20) // 4
        cur, cur_n = [], 0
        for p in reversed(self.params):
            if cur and cur_n + p.numel() > cap:
                self._make_bucket(cur)
                cur, cur_n = [], 0
            cur.append(p)
            cur_n += p.numel()
        if cur:
            self._make_bucket(cur)
        self.cuda = self.params[0].is_cuda
        self.comm_stream = torch.cuda.Stream() if self.cuda else None
        for bi, (_, ps) in enumerate(self.buckets):
            for p in ps:
                p.register_post_accumulate_grad_hook(self._make_hook(bi))
        self.zero_grad()

    def _make_bucket(self, ps):
        n = sum(p.numel() for p in ps)
        flat = torch.zeros(n, dtype=torch.float32, device=ps[0].device)
        off = 0
        for p in ps:
            p.grad = flat[off:off + p.numel()].view_as(p)
            off += p.numel()
        self.buckets.append((flat, list(ps)))

    def _make_hook(self, bi):
        def hook(param):
            self._pending[bi] -= 1
            if self._pending[bi] == 0:
                self._launch(bi)
        return hook

    def zero_grad(self):
        self._works = []
        for bi, (flat, ps) in enumerate(self.buckets):
            flat.zero_()
            self._pending[bi] = len(ps)
            off = 0
            for p in ps:  # re-attach views if an optimizer/zero_grad(set_to_none=True) dropped them
                if p.grad is None or p.grad.data_ptr() != flat.data_ptr() + 4 * off:
                    p.grad = flat[off:off + p.numel()].view_as(p)
                off += p.numel()

    def _launch(self, bi):
        if self.world == 1:
            return
        flat = self.buckets[bi][0]
        if self.cuda:
            self.comm_stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.comm_stream):
                flat.mul_(1.0 / self.world)
                self._works.append(dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True))
        else:
            flat.mul_(1.0 / self.world)
            self._works.append(dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True))

    def wait(self):
        """Call after backward(): launches any bucket whose hook count did not reach zero (unused parameters) and makes
        the compute stream wait for all exchanges."""
        for bi in range(len(self.buckets)):
            if self._pending[bi] > 0:
                self._pending[bi] = 0
                self._launch(bi)
        for w in self._works:
            w.wait()
        if self.cuda and self.world > 1:
            torch.cuda.current_stream().wait_stream(self.comm_stream)
        self._works = []

    def payload_bytes(self):
        return sum(f.numel() * 4 for f, _ in self.buckets)
