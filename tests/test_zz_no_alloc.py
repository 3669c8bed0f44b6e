"""-m gpu, collected LAST on purpose (a failure here must not hide other tests).

`ctr_reduce_static` -- the allocation-free variant of `ctr_reduce` (include/ctr_b200.h; no counterpart in the
reference, added for callers that cannot call cudaMalloc).  Its block partials
live in a file-scope `static __device__` array inside libctr_b200.so.  Checked through the C-ABI against the existing
entry point on two calls in a row (the second smaller than the first, so stale partials of the first call are sitting
in the array), for the three modes; the kernels and block slices are the same, so the bar is bit equality, which is
inside every tolerance the other tests use.  Passes alone in a fresh process and after the rest of the suite."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def hd():
    from deep_learning_b200 import _lib
    return _lib.Handle(0)


def _st():
    return torch.cuda.current_stream().cuda_stream


def _both(hd, x, mode):
    from deep_learning_b200._lib import ptr
    n = x.numel()
    ws = torch.zeros(8192, dtype=torch.uint8, device="cuda")
    want = torch.full((1,), float("nan"), device="cuda")
    got = torch.full((1,), float("nan"), device="cuda")
    hd.call("ctr_reduce", _st(), ptr(x), n, mode, ptr(want), ptr(ws), ws.numel())
    hd.call("ctr_reduce_static", _st(), ptr(x), n, mode, ptr(got))
    torch.cuda.synchronize()
    return float(got.item()), float(want.item())


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_reduce_static_equals_reduce_on_two_calls_in_a_row(hd, mode):
    g = torch.Generator(device="cuda")
    g.manual_seed(100 + mode)
    first = torch.randn(5_000_011, generator=g, device="cuda")          # 1024 partials: fills the whole static array
    second = torch.randn(70_001, generator=g, device="cuda") * 3.0      # 18 partials: the rest of the array is stale
    f = {0: lambda a: a, 1: lambda a: a * a, 2: lambda a: a.abs()}[mode]
    for x in (first, second):
        got, want = _both(hd, x, mode)
        assert got == want, (mode, x.numel(), got, want)                 # same kernels, same slices: bit-identical
        ref = float(f(x.double()).sum().item())
        assert abs(got - ref) <= 1e-5 * max(1.0, float(f(x.double()).abs().sum().item())), (mode, got, ref)


def test_reduce_static_allocates_nothing_and_rejects_oversize(hd):
    from deep_learning_b200 import _lib
    from deep_learning_b200._lib import ptr
    x = torch.ones(4096 * 3 + 5, device="cuda")
    out = torch.zeros(1, device="cuda")
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    reserved0 = torch.cuda.memory_reserved()
    for _ in range(3):
        hd.call("ctr_reduce_static", _st(), ptr(x), x.numel(), 0, ptr(out))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert float(out.item()) == float(x.numel())
    assert torch.cuda.memory_reserved() == reserved0 and free1 == free0          # neither torch nor the driver handed out memory
    # beyond the stated maximum: a clear error instead of a silent wrong answer (no memory of that size is touched)
    with pytest.raises(_lib.CtrError, match="exceeds the static scratch's limit"):
        hd.call("ctr_reduce_static", _st(), ptr(x), (1 << 28) + 1, 0, ptr(out))
    # n = 0 is a valid empty sum
    hd.call("ctr_reduce_static", _st(), ptr(x), 0, 1, ptr(out))
    torch.cuda.synchronize()
    assert float(out.item()) == 0.0
