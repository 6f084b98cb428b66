"""l2hmc_col_sum_static: the no-allocation variant of l2hmc_col_sum (scratch in a file-scope device array of the
library).  It has no counterpart in the reference; it exists for callers that cannot call cudaMalloc.  Checked through
the C ABI against l2hmc_col_sum (same kernels, same summation order: identical bits) on two calls in a row that share
the built-in scratch, and against a float64 sum.  Collected last on purpose."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_FLOATS = 1 << 20   # L2HMC_COL_SUM_STATIC_MAX_FLOATS (include/l2hmc_b200.h)


def _slabs(rows):
    return max(1, min(512, rows // 32))


def _both(lib, _lib, B, rows, cols, ld):
    """(l2hmc_col_sum with a caller workspace, l2hmc_col_sum_static) on the same input."""
    need = int(lib.l2hmc_col_sum_workspace_bytes(rows, cols))
    ws = torch.empty(max(need, 4), dtype=torch.uint8, device="cuda")
    ref = torch.full((cols,), float("nan"), device="cuda")
    got = torch.full((cols,), float("nan"), device="cuda")
    _lib.check(lib.l2hmc_col_sum(_lib.dptr(B), ld, rows, cols, _lib.dptr(ref), ws.data_ptr(), need, _lib.stream_ptr()))
    _lib.check(lib.l2hmc_col_sum_static(_lib.dptr(B), ld, rows, cols, _lib.dptr(got), _lib.stream_ptr()))
    torch.cuda.synchronize()
    return ref.cpu().numpy(), got.cpu().numpy()


def test_col_sum_static_equals_col_sum_on_consecutive_calls():
    from l2hmc_b200 import _lib
    lib = _lib.get_lib()
    gen = torch.Generator(device="cpu").manual_seed(7)
    # a large call first, then a smaller one whose partials sit on top of the first call's leftovers, then a view with
    # a row stride (ld > cols) and a single-slab case
    for rows, cols, ld in ((40960, 1536, 1536), (777, 130, 130), (4096, 384, 1536), (31, 5, 5)):
        assert _slabs(rows) * cols <= MAX_FLOATS
        B = torch.randn(rows, ld, generator=gen).cuda()
        ref, got = _both(lib, _lib, B, rows, cols, ld)
        assert np.array_equal(ref, got), f"static scratch differs from the workspace path at {rows}x{cols}"
        exact = B[:, :cols].double().sum(0).cpu().numpy()
        # fp32 sums of `rows` unit normals: tolerance = a few ulps of the partial sums' scale
        np.testing.assert_allclose(got, exact, rtol=0, atol=2e-6 * rows ** 0.5 * 8)
    # the same input twice in a row: identical bits both times (nothing stale is read)
    B = torch.randn(2048, 512, generator=gen).cuda()
    a = _both(lib, _lib, B, 2048, 512, 512)[1]
    b = _both(lib, _lib, B, 2048, 512, 512)[1]
    assert np.array_equal(a, b)


def test_col_sum_static_rejects_sizes_beyond_its_scratch():
    from l2hmc_b200 import _lib
    lib = _lib.get_lib()
    rows, cols = 1 << 15, 4096                      # 512 slabs x 4096 columns = 2 Mi floats > 1 Mi
    assert _slabs(rows) * cols > MAX_FLOATS
    B = torch.zeros(8, device="cuda")               # never read: the call is refused before any launch
    out = torch.zeros(8, device="cuda")
    rc = lib.l2hmc_col_sum_static(_lib.dptr(B), cols, rows, cols, _lib.dptr(out), _lib.stream_ptr())
    assert rc == _lib.L2HMC_ERR_INVALID
    msg = lib.l2hmc_last_error().decode()
    assert "L2HMC_COL_SUM_STATIC_MAX_FLOATS" in msg and "l2hmc_col_sum" in msg
    # the largest size that fits still runs
    rows, cols = 1 << 14, 2048                      # 512 x 2048 = 1 Mi floats exactly
    assert _slabs(rows) * cols == MAX_FLOATS
    Bm = torch.ones(rows, cols, device="cuda")
    ref, got = _both(lib, _lib, Bm, rows, cols, cols)
    assert np.array_equal(ref, got) and np.all(got == rows)
