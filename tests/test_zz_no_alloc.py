"""The allocation-free entry point f5b_op_sumsq_static (no counterpart in the reference; added
for callers that cannot call cudaMalloc): its scratch is a file-scope `static __device__` array of the
library.  Checked through the C ABI against the existing entry point f5b_op_sumsq (caller-supplied scratch) on two calls
in a row -- the second with other data and another length, so stale partials of the first call would show -- and
against a float64 sum.  Collected last on purpose."""
import pytest
import torch

import helpers  # noqa: F401  (sys.path)

pytestmark = pytest.mark.gpu

REL_TOL = 1e-12  # both entry points run the same fixed reduction tree in double precision


def _both(L, _lib, x, dev):
    st = _lib.current_stream_ptr(dev)
    idx = _lib.device_index(dev)
    scratch = torch.empty(1024, device=dev, dtype=torch.float64)
    ref = torch.full((1,), -1.0, device=dev, dtype=torch.float64)
    got = torch.full((1,), -2.0, device=dev, dtype=torch.float64)
    _lib.check(L.f5b_op_sumsq(x.data_ptr(), x.numel(), scratch.data_ptr(), ref.data_ptr(), idx, st))
    _lib.check(L.f5b_op_sumsq_static(x.data_ptr(), x.numel(), got.data_ptr(), idx, st))
    torch.cuda.synchronize(dev)
    return float(ref), float(got)


def test_sumsq_static_equals_sumsq_on_two_calls_in_a_row():
    from f5b200 import _lib

    L = _lib.lib()
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(5)
    first = (torch.randn(3_000_017, generator=g) * 3.0).to(dev)   # every block has work
    second = torch.randn(777, generator=g).to(dev)                # most blocks have none: their partials must read 0
    for x in (first, second):
        ref, got = _both(L, _lib, x, dev)
        exact = float((x.double() ** 2).sum())
        assert abs(got - ref) <= REL_TOL * abs(ref), (got, ref)
        assert abs(got - exact) <= 1e-9 * exact, (got, exact)
    # the same input again after the short one: nothing of the previous call leaks into the result
    ref, got = _both(L, _lib, first, dev)
    assert abs(got - ref) <= REL_TOL * abs(ref)


def test_sumsq_static_rejects_sizes_beyond_its_maximum():
    from f5b200 import _lib

    L = _lib.lib()
    dev = torch.device("cuda", 0)
    x = torch.ones(16, device=dev)
    out = torch.zeros(1, device=dev, dtype=torch.float64)
    rc = L.f5b_op_sumsq_static(x.data_ptr(), (1 << 31) + 1, out.data_ptr(), 0, _lib.current_stream_ptr(dev))
    assert rc == _lib.ERR_INVALID
    assert b"exceeds the maximum" in L.f5b_last_error()
    torch.cuda.synchronize(dev)
    assert float(out) == 0.0  # rejected before anything was launched
