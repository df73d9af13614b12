"""m3ae_skinny_linear_static — the allocation-free C-ABI entry point (include/m3ae_b200.h; no counterpart in the reference,
added for callers that cannot call cudaMalloc) — against the existing m3ae_skinny_linear on the same inputs, two calls in
a row through the shared static scratch.  The file name keeps it last in pytest's collection order."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

import m3ae_b200  # noqa: E402,F401
from m3ae_b200 import _lib as L  # noqa: E402
from m3ae_b200 import ops  # noqa: E402

DEV = "cuda"
TOL = 4e-3  # the bf16 gate of tests/test_kernels_gpu.py::test_skinny_linear_matches_layernorm_plus_gemm


def rel_err(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


def rnd(*shape, seed=0, scale=1.0):
    g = torch.Generator(device="cpu").manual_seed(seed + sum(shape))
    return (torch.randn(*shape, generator=g) * scale).to(DEV)


def run_static(launch):
    """The same cached parameter block through the new entry point, with the caller's scratch pointers cleared."""
    p = L.SkinnyParams()
    C.memmove(C.byref(p), C.byref(launch.p), C.sizeof(p))
    p.ws, p.counters = None, None
    L.check(L.load().m3ae_skinny_linear_static(C.byref(p), ops._stream()), "m3ae_skinny_linear_static")


@pytest.mark.parametrize("M", [128, 37])
def test_static_scratch_split_k_equals_caller_scratch_twice_in_a_row(M):
    D, FF = 768, 3072
    bf = torch.bfloat16
    w2, b2 = (0.05 * rnd(D, FF, seed=11)).to(bf), rnd(D, seed=12)
    ws, ctr = torch.zeros(128 * D, device=DEV), torch.zeros(256, dtype=torch.int32, device=DEV)
    y_ref, y = torch.empty(M, D, dtype=bf, device=DEV), torch.empty(M, D, dtype=bf, device=DEV)
    inputs = [(rnd(M, FF, seed=20 + i).to(bf), rnd(M, D, seed=30 + i).to(bf)) for i in range(2)]
    free_before = None
    for call, (h, res) in enumerate(inputs):  # different inputs per call: leftovers of call 0 in the scratch would show in call 1
        ops.SkinnyLinear(h, w2, y_ref, bias=b2, residual=res, workspace=ws, counters=ctr).run()
        launch = ops.SkinnyLinear(h, w2, y, bias=b2, residual=res, workspace=ws, counters=ctr)
        assert launch.p.ksplit > 1, "K = 3072 must take the split-K path (the one that needs scratch)"
        y.fill_(float("nan"))
        torch.cuda.synchronize()
        if call == 1:
            free_before = torch.cuda.mem_get_info()[0]
        run_static(launch)
        torch.cuda.synchronize()
        if call == 1:
            assert torch.cuda.mem_get_info()[0] == free_before, "the static-scratch call must not allocate device memory"
        e = rel_err(y, y_ref)
        assert torch.isfinite(y.float()).all() and e < TOL, f"call {call}, M={M}: {e:.3e}"
    assert (ws == 0).all() and (ctr == 0).all(), "the caller's scratch is not touched by the static variant"


def test_static_scratch_unsplit_projection_and_size_limit():
    M, D = 64, 768
    bf = torch.bfloat16
    x, w, b = rnd(M, D, seed=1).to(bf), (0.05 * rnd(D, D, seed=2)).to(bf), rnd(D, seed=3)
    ref, out = torch.empty(M, D, dtype=bf, device=DEV), torch.empty(M, D, dtype=bf, device=DEV)
    launch = ops.SkinnyLinear(x, w, ref, bias=b)
    launch.run()
    run_static(ops.SkinnyLinear(x, w, out, bias=b))
    assert torch.equal(out, ref), "no split-K: the same kernel on the same operands, bit for bit"
    # beyond the stated maximum (1024 output columns): a clear error, nothing launched
    n_big = 1056
    wb, ob = (0.05 * rnd(n_big, D, seed=4)).to(bf), torch.zeros(M, n_big, dtype=bf, device=DEV)
    with pytest.raises(RuntimeError, match="exceed the static scratch"):
        run_static(ops.SkinnyLinear(x, wb, ob))
    torch.cuda.synchronize()
    assert (ob == 0).all()
