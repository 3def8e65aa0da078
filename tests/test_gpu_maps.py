"""LinearMapImpl::Apply parity (reference tests: src/epsilon/linear/dense_matrix_impl_test.cc:24-29,
linear_map_test.cc, kronecker_product_impl_test.cc:9-20; tolerance 1e-8 there, tighter here)."""
import numpy as np
import pytest
import scipy.sparse as sp

from epsilon_b200.problem import LinearMapSpec
from epsilon_b200.solver import LinearMap

pytestmark = pytest.mark.gpu


def _check(ctx, spec, rng, tol=1e-12):
    lm = LinearMap(ctx, spec)
    m, n = lm.shape
    assert (m, n) == (spec.m, spec.n)
    x = rng.standard_normal(n)
    y = lm.apply(x)
    ref = spec.apply(x)
    scale = max(1.0, np.abs(ref).max())
    assert np.abs(y - ref).max() <= tol * scale * max(1, n) ** 0.5
    yt = rng.standard_normal(m)
    t = lm.transpose()
    assert t.shape == (n, m)
    xt = t.apply(yt)
    reft = np.asarray(spec.to_scipy().T @ yt).ravel()
    scale = max(1.0, np.abs(reft).max())
    assert np.abs(xt - reft).max() <= tol * scale * max(1, m) ** 0.5


@pytest.mark.parametrize("m,n", [(1, 1), (3, 5), (64, 64), (1000, 37), (37, 1000), (513, 2049), (4096, 300)])
def test_dense(ctx, m, n):
    rng = np.random.default_rng(m * 1000 + n)
    _check(ctx, LinearMapSpec.Dense(rng.standard_normal((m, n))), rng)


@pytest.mark.parametrize("m,n,d", [(5, 7, 0.5), (100, 50, 0.1), (2000, 300, 0.01), (300, 2000, 0.05),
                                   (1000, 1000, 0.3), (50, 40, 0.0)])
def test_sparse(ctx, m, n, d):
    rng = np.random.default_rng(m + n)
    a = sp.random(m, n, density=d, random_state=np.random.RandomState(1), format="csc")
    _check(ctx, LinearMapSpec.Sparse(a), rng)


def test_sparse_ragged_rows(ctx):
    rng = np.random.default_rng(3)
    a = sp.lil_matrix((200, 300))
    a[0, :] = rng.standard_normal(300)          # one full row
    a[5, 3] = 2.0                               # one singleton
    a[100:150, 10] = rng.standard_normal((50, 1))  # one full column segment; many empty rows
    _check(ctx, LinearMapSpec.Sparse(a.tocsc()), rng)


def test_diagonal_scalar(ctx):
    rng = np.random.default_rng(0)
    _check(ctx, LinearMapSpec.Diagonal(rng.standard_normal(777)), rng)
    _check(ctx, LinearMapSpec.Scalar(-2.5, 1234), rng)


def test_kronecker(ctx):
    rng = np.random.default_rng(1)
    a = LinearMapSpec.Dense(rng.standard_normal((3, 4)))
    b = LinearMapSpec.Sparse(sp.random(5, 6, density=0.5, random_state=np.random.RandomState(2)))
    _check(ctx, LinearMapSpec.Kronecker(a, b), rng)


@pytest.mark.parametrize("m,n,d", [(3000, 2000, 0.05), (70000, 30001, 0.0005), (301, 60000, 0.02),
                                   (60001, 257, 0.02)])
def test_sparse_sell(ctx, m, n, d, monkeypatch):
    """Column-tiled SELL path (large matrices; forced here for small ones as well)."""
    monkeypatch.setenv("EPS_SPARSE_FORMAT", "sell")
    rng = np.random.default_rng(m + n)
    a = sp.random(m, n, density=d, random_state=np.random.RandomState(5), format="csc")
    a.data = rng.standard_normal(a.nnz)
    _check(ctx, LinearMapSpec.Sparse(a), rng)


def test_sparse_sell_ragged(ctx, monkeypatch):
    monkeypatch.setenv("EPS_SPARSE_FORMAT", "sell")
    rng = np.random.default_rng(11)
    a = sp.lil_matrix((500, 70000))
    a[0, :] = rng.standard_normal(70000)            # a row longer than one tile
    a[7, 69999] = 3.0
    a[100:400, 25600] = rng.standard_normal((300, 1))  # first column of the second tile
    _check(ctx, LinearMapSpec.Sparse(a.tocsc()), rng)


@pytest.mark.parametrize("m,n,d", [(5, 7, 0.5), (31, 33, 0.2), (1, 40000, 0.01), (40000, 1, 0.3), (64, 64, 0.0),
                                   (1025, 51201, 0.003)])
def test_sparse_sell_small_shapes(ctx, m, n, d, monkeypatch):
    """Stream layout edge cases: fewer rows than a slice, fewer slices than warps, a single row or
    column, no stored entries at all, a tile boundary that is not a multiple of the tile width."""
    monkeypatch.setenv("EPS_SPARSE_FORMAT", "sell")
    rng = np.random.default_rng(m * 7 + n)
    a = sp.random(m, n, density=d, random_state=np.random.RandomState(9), format="csc")
    a.data = rng.standard_normal(a.nnz)
    _check(ctx, LinearMapSpec.Sparse(a), rng)


def test_sparse_sell_skewed_rows(ctx, monkeypatch):
    """Row lengths from 0 to several thousand inside one visit (sorting, empty trailing slices)."""
    monkeypatch.setenv("EPS_SPARSE_FORMAT", "sell")
    rng = np.random.default_rng(21)
    m, n = 5000, 30000
    lens = np.minimum((rng.pareto(1.2, m) * 3).astype(int), 4000)
    lens[::7] = 0
    rows = np.repeat(np.arange(m), lens)
    cols = np.concatenate([rng.choice(n, k, replace=False) for k in lens if k]) if lens.sum() else np.zeros(0, int)
    a = sp.csc_matrix((rng.standard_normal(rows.size), (rows, cols)), shape=(m, n))
    _check(ctx, LinearMapSpec.Sparse(a), rng)


def test_buffers_may_outlive_their_context():
    """A host language finalises objects in any order: maps and solvers destroyed after their
    context must free their device memory without touching it (stream-ordered pool, ctx.cu)."""
    from epsilon_b200 import problems
    from epsilon_b200.solver import Context, ProxADMMSolver
    c = Context(0)
    rng = np.random.default_rng(3)
    a = sp.random(4000, 3000, density=0.03, random_state=np.random.RandomState(4), format="csc")
    lm = LinearMap(c, LinearMapSpec.Sparse(a))
    assert np.isfinite(lm.apply(rng.standard_normal(3000))).all()
    prob, _ = problems.lasso(200, 50, seed=1)
    s = ProxADMMSolver(prob, {}, ctx=c)
    s.solve()
    c.close()
    lm.close()
    s.close()
    c2 = Context(0)                       # the device is still usable
    lm2 = LinearMap(c2, LinearMapSpec.Sparse(a))
    assert np.isfinite(lm2.apply(rng.standard_normal(3000))).all()
    lm2.close()
    c2.close()


def _raw_csc(ctx, m, n, indptr, indices, data):
    """eps_map_csc on raw arrays (no scipy canonicalisation in between)."""
    import ctypes as C
    from epsilon_b200 import capi
    indptr = np.ascontiguousarray(indptr, dtype=np.int32)
    indices = np.ascontiguousarray(indices, dtype=np.int32)
    data = np.ascontiguousarray(data, dtype=np.float64)
    h = C.c_void_p()
    rc = ctx.lib.eps_map_csc(ctx.handle, m, n, len(data), capi.iptr(indptr), capi.iptr(indices),
                             capi.dptr(data), C.byref(h))
    return rc, h


@pytest.mark.parametrize("defect", ["row_out_of_range", "negative_row", "colptr_not_monotone"])
def test_malformed_csc_is_an_error_code_not_a_device_fault(ctx, defect):
    """a malformed wire matrix must come back as EPS_ERR_INVALID; the context stays usable."""
    rng = np.random.default_rng(5)
    a = sp.random(300, 40, density=0.1, random_state=7, format="csc")
    indptr, indices, data = a.indptr.copy(), a.indices.copy(), a.data.copy()
    if defect == "row_out_of_range":
        indices[len(indices) // 2] = 300
    elif defect == "negative_row":
        indices[3] = -1
    else:
        indptr[10], indptr[11] = indptr[11], indptr[10]
        assert indptr[10] > indptr[11]
    rc, h = _raw_csc(ctx, 300, 40, indptr, indices, data)
    assert rc == -1 and not h.value
    assert b"malformed CSC" in ctx.lib.eps_last_error()
    # no sticky CUDA error: the same context still computes
    lm = LinearMap(ctx, LinearMapSpec.Sparse(a))
    x = rng.standard_normal(40)
    assert np.abs(lm.apply(x) - a @ x).max() <= 1e-13


@pytest.mark.parametrize("fmt", ["csr", "sell"])
def test_unsorted_csc_columns_are_sorted_at_ingest(ctx, fmt, monkeypatch):
    """Row indices in arbitrary order inside a column are legal CSC (Eigen accepts them); the device kernels
    need them ascending, so ingest sorts them."""
    monkeypatch.setenv("EPS_SPARSE_FORMAT", fmt)
    rng = np.random.default_rng(6)
    a = sp.random(5000, 700, density=0.02, random_state=8, format="csc")
    a.sort_indices()
    indptr, indices, data = a.indptr.copy(), a.indices.copy(), a.data.copy()
    for j in range(700):
        lo, hi = indptr[j], indptr[j + 1]
        perm = rng.permutation(hi - lo)
        indices[lo:hi] = indices[lo:hi][perm]
        data[lo:hi] = data[lo:hi][perm]
    rc, h = _raw_csc(ctx, 5000, 700, indptr, indices, data)
    assert rc == 0
    lm = LinearMap(ctx, _handle=h)
    x, y = rng.standard_normal(700), rng.standard_normal(5000)
    assert np.abs(lm.apply(x) - a @ x).max() <= 1e-13 * 10
    assert np.abs(lm.transpose().apply(y) - a.T @ y).max() <= 1e-13 * 100


def test_inverse_per_map_type(ctx):
    """LinearMapImpl::Inverse (linear_map.h:49; block_matrix_test.cc:78-93 checks block-diagonal inverses
    against Eigen::LLT at 1e-8)."""
    from epsilon_b200.capi import EpsError
    rng = np.random.default_rng(12)
    n = 150
    x = rng.standard_normal(n)
    # scalar, diagonal with a zero entry (kept zero, diagonal_matrix_impl.cc:15-23)
    assert np.allclose(LinearMap(ctx, LinearMapSpec.Scalar(-4.0, n)).inverse().apply(x), x / -4.0, rtol=1e-15)
    d = rng.standard_normal(n)
    d[7] = 0.0
    ref = np.where(d != 0, x / np.where(d != 0, d, 1.0), 0.0)
    assert np.allclose(LinearMap(ctx, LinearMapSpec.Diagonal(d)).inverse().apply(x), ref, rtol=1e-15)
    # dense symmetric positive definite, and negative definite (the -I - 2a AA' pivots of BlockCholesky)
    b = rng.standard_normal((n, 40))
    spd = b @ b.T + np.eye(n)
    for mat in (spd, -spd):
        inv = LinearMap(ctx, LinearMapSpec.Dense(mat)).inverse()
        assert inv.shape == (n, n)
        assert np.abs(inv.apply(x) - np.linalg.solve(mat, x)).max() <= 1e-10 * np.abs(x).max()
    # sparse: scalar matrix -> scalar inverse; general -> dense inverse (sparse_matrix_impl.cc:60-78)
    assert np.allclose(LinearMap(ctx, LinearMapSpec.Sparse(2.5 * sp.identity(n, format="csc"))).inverse().apply(x), x / 2.5)
    s = sp.random(n, n, density=0.05, random_state=3, format="csc")
    ssp = (s @ s.T + sp.identity(n)).tocsc()
    got = LinearMap(ctx, LinearMapSpec.Sparse(ssp)).inverse().apply(x)
    assert np.abs(got - np.linalg.solve(ssp.toarray(), x)).max() <= 1e-10 * np.abs(x).max()
    # Kronecker: inv(A) (x) inv(B)
    a2 = rng.standard_normal((4, 6))
    a2 = a2 @ a2.T + np.eye(4)
    k = LinearMapSpec.Kronecker(LinearMapSpec.Dense(a2), LinearMapSpec.Scalar(0.5, 10))
    xk = rng.standard_normal(40)
    got = LinearMap(ctx, k).inverse().apply(xk)
    assert np.abs(got - np.linalg.solve(np.kron(a2, 0.5 * np.eye(10)), xk)).max() <= 1e-11
    # indefinite: refused, not approximated
    ind = np.diag(np.concatenate([np.ones(5), -np.ones(5)]))
    with pytest.raises(EpsError):
        LinearMap(ctx, LinearMapSpec.Dense(ind)).inverse()
    # the context survives the refused factorisation
    assert np.allclose(LinearMap(ctx, LinearMapSpec.Scalar(2.0, 10)).inverse().apply(np.ones(10)), 0.5)
