"""world_size-2 gloo test (CPU) of the gradient exchange used for N>1 GPUs: averaged gradients must equal the
single-process gradient of the concatenated batch."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _model():
    torch.manual_seed(0)
    return torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.Tanh(), torch.nn.Linear(32, 4), torch.nn.Linear(4, 4))


def _worker(rank, world, port, q, overlap):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, ROOT)
    import importlib
    ddp = importlib.import_module("mm-vqa-healthcare_b200.ddp")
    m = _model()
    red = ddp.GradientAllReducer(list(m.parameters()), world, bucket_mb=0, overlap=overlap)  # bucket_mb=0 -> one bucket per parameter
    g = torch.Generator().manual_seed(5)
    x = torch.randn(8, 16, generator=g)
    for it in range(2):  # second iteration checks zero_grad / hook re-arming
        red.zero_grad()
        xs = x[rank * 4:(rank + 1) * 4]
        m(xs).square().sum().backward()
        red.wait()
    q.put((rank, [p.grad.numpy().copy() for p in m.parameters()]))  # by value: fd-shared tensors die with the worker
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("overlap", [True, False])
def test_two_rank_gradient_average_matches_single_process(overlap):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000 + (0 if overlap else 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q, overlap)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    m = _model()
    g = torch.Generator().manual_seed(5)
    x = torch.randn(8, 16, generator=g)
    (m(x).square().sum() / 2).backward()  # mean over 2 ranks of per-rank sums
    for r in range(2):
        for a, p in zip(got[r], m.parameters()):
            assert torch.allclose(torch.from_numpy(a), p.grad, rtol=1e-5, atol=1e-6)


class _FlatGradLinear(torch.autograd.Function):
    """y = x W1^T W2^T whose backward returns dW1, dW2 as views of ONE flat buffer — the way the fused BertCrossLayer
    node hands out a layer's 26 parameter gradients (functional.py::_backward_streams)."""

    @staticmethod
    def forward(ctx, x, w1, w2):
        h = x @ w1.t()
        ctx.save_for_backward(x, w1, w2, h)
        return h @ w2.t()

    @staticmethod
    def backward(ctx, dy):
        x, w1, w2, h = ctx.saved_tensors
        flat = torch.zeros(w1.numel() + w2.numel())
        g1, g2 = flat[:w1.numel()].view_as(w1), flat[w1.numel():].view_as(w2)
        g2.copy_(dy.t() @ h)
        dh = dy @ w2
        g1.copy_(dh.t() @ x)
        return dh @ w1, g1, g2


class _FlatModel(torch.nn.Module):
    def __init__(self):
        super().__init__()
        torch.manual_seed(0)
        self.w1 = torch.nn.Parameter(torch.randn(12, 16) * 0.1)
        self.w2 = torch.nn.Parameter(torch.randn(6, 12) * 0.1)
        self.head = torch.nn.Linear(6, 3)

    def forward(self, x):
        return self.head(_FlatGradLinear.apply(x, self.w1, self.w2))


def _worker_flat(rank, world, port, q, overlap):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, ROOT)
    import importlib
    ddp = importlib.import_module("mm-vqa-healthcare_b200.ddp")
    m = _FlatModel()
    red = ddp.GradientAllReducer(list(m.parameters()), world, overlap=overlap)
    g = torch.Generator().manual_seed(5)
    x = torch.randn(8, 16, generator=g)
    info = {}
    for it in range(3):  # step 0 discovers the group, steps 1-2 exchange it from the hooks
        red.zero_grad()
        if it > 0:
            assert m.w1.grad is None and m.w2.grad is None and m.head.weight.grad is not None
        m(x[rank * 4:(rank + 1) * 4]).square().sum().backward()
        red.wait()
    info["groups"] = [len(ps) for ps in red.groups]
    info["bucketed"] = sum(len(ps) for _, ps in red.buckets)
    info["shared"] = m.w1.grad.untyped_storage().data_ptr() == m.w2.grad.untyped_storage().data_ptr()
    q.put((rank, [p.grad.numpy().copy() for p in m.parameters()], info))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("overlap", [True, False])
def test_zero_copy_group_exchanges_a_layers_flat_gradient_buffer_in_place(overlap):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 33500 + os.getpid() % 2000 + (0 if overlap else 2000)
    procs = [ctx.Process(target=_worker_flat, args=(r, 2, port, q, overlap)) for r in range(2)]
    for p in procs:
        p.start()
    got = {r: (gs, info) for r, gs, info in (q.get(timeout=120) for _ in range(2))}
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    m = _FlatModel()
    g = torch.Generator().manual_seed(5)
    x = torch.randn(8, 16, generator=g)
    (m(x).square().sum() / 2).backward()
    for r in range(2):
        gs, info = got[r]
        assert info["groups"] == [2] and info["bucketed"] == 2 and info["shared"], info  # w1|w2 zero-copy, head bucketed
        for a, p in zip(gs, m.parameters()):
            assert torch.allclose(torch.from_numpy(a), p.grad, rtol=1e-5, atol=1e-6)
