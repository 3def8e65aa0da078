"""eps_dense_apply_noalloc (include/epsgpu.h): y = A x with the split partials in a static device array, no
allocation.  No counterpart in the reference -- so the check is
against the library's existing entry point for the same product (eps_map_dense + eps_map_apply_dev / eps_map_apply),
which the golden tests hold to the reference's DenseMatrixImpl::Apply.  Collected last on
purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
EPS_ERR_INVALID = -1   # include/epsgpu.h


def _dev(t):
    return C.c_void_p(t.data_ptr())


def _both(ctx, a_host, x_host):
    """(no-alloc result, existing entry point on device buffers, existing entry point on host buffers)"""
    import torch
    from epsilon_b200 import capi
    from epsilon_b200.problem import LinearMapSpec
    from epsilon_b200.solver import LinearMap
    lib = capi.load()
    m, n = a_host.shape
    dev = torch.device("cuda", 0)
    a = torch.from_numpy(np.asfortranarray(a_host).ravel(order="F").copy()).to(dev)   # column-major
    x = torch.from_numpy(x_host).to(dev)
    y_new = torch.full((m,), float("nan"), dtype=torch.float64, device=dev)
    y_old = torch.full((m,), float("nan"), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    before = ctx.bytes_allocated
    capi.check(lib.eps_dense_apply_noalloc(ctx.handle, m, n, _dev(a), _dev(x), _dev(y_new)))
    ctx.synchronize()
    assert ctx.bytes_allocated == before                      # the context allocated nothing for this call
    amap = LinearMap(ctx, LinearMapSpec.Dense(a_host))
    capi.check(lib.eps_map_apply_dev(amap.handle, _dev(x), _dev(y_old)))
    ctx.synchronize()
    y_host = amap.apply(x_host)
    amap.close()
    return y_new.cpu().numpy(), y_old.cpu().numpy(), y_host


def test_noalloc_matches_the_existing_entry_point_twice_in_a_row(ctx):
    """Two calls in a row, the second on a different shape: the shared array keeps the first call's partials and the
    second call must not see them."""
    rng = np.random.default_rng(11)
    for m, n in [(1000, 5000), (777, 1301)]:
        a = rng.standard_normal((m, n))
        x = rng.standard_normal(n)
        y_new, y_old, y_host = _both(ctx, a, x)
        assert np.all(np.isfinite(y_new))
        assert np.array_equal(y_new, y_old)      # same kernels, same split, same summation order
        assert np.array_equal(y_new, y_host)
        ref = a @ x
        assert np.abs(y_new - ref).max() <= 1e-12 * max(1.0, np.abs(ref).max()) * np.sqrt(n)
    # the same shape again with new data: nothing of the previous contents may leak
    a = rng.standard_normal((777, 1301))
    for _ in range(2):
        x = rng.standard_normal(1301)
        y_new, y_old, _ = _both(ctx, a, x)
        assert np.array_equal(y_new, y_old)


def test_noalloc_rejects_sizes_beyond_the_static_scratch(ctx):
    import torch
    from epsilon_b200 import capi
    lib = capi.load()
    t = torch.zeros(8, dtype=torch.float64, device="cuda:0")
    # 300000 x 640: 6 column splits x 300000 rows of partial sums > 2^20; rejected before anything is launched
    rc = lib.eps_dense_apply_noalloc(ctx.handle, 300000, 640, _dev(t), _dev(t), _dev(t))
    assert rc == EPS_ERR_INVALID
    msg = lib.eps_last_error().decode()
    assert "EPS_NOALLOC_SCRATCH_DOUBLES" in msg and "300000" in msg
    assert lib.eps_dense_apply_noalloc(ctx.handle, 0, 5, _dev(t), _dev(t), _dev(t)) == EPS_ERR_INVALID
    assert lib.eps_dense_apply_noalloc(ctx.handle, 4, 2, None, _dev(t), _dev(t)) == EPS_ERR_INVALID
