"""world_size-2 gloo test (CPU) of the gradient exchange used for N>1 GPUs: averaged gradients must equal the
single-process gradient of the concatenated batch."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _model():
    torch.manual_seed(0)
    return torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.Tanh(), torch.nn.Linear(32, 4), torch.nn.Linear(4, 4))


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, ROOT)
    import importlib
    ddp = importlib.import_module("mm-vqa-healthcare_b200.ddp")
    m = _model()
    red = ddp.GradientAllReducer(list(m.parameters()), world, bucket_mb=0)  # bucket_mb=0 -> one bucket per parameter
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


def test_two_rank_gradient_average_matches_single_process():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
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
