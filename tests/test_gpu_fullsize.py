"""Parity at BASELINE.json's full size (sparse lasso 1M x 100k, nnz 1e8), where the compiled
reference cannot run (it would form and invert a dense 100k x 100k matrix,
src/epsilon/linear/sparse_matrix_impl.cc:60-78): the linear maps are checked against scipy on
the same bytes and through size-independent properties (adjointness, linearity), the whole
solve through the agreement of two independent device paths (stream-SELL kernels + fused CG
update vs plain CSR kernels + unfused CG) and the optimality conditions of the lasso."""
import numpy as np
import pytest

from epsilon_b200 import problems
from epsilon_b200.problem import LinearMapSpec
from epsilon_b200.solver import LinearMap, ProxADMMSolver

pytestmark = pytest.mark.gpu

M, N, DENSITY = 1_000_000, 100_000, 1e-3


@pytest.fixture(scope="module")
def full():
    prob, info = problems.lasso_sparse_panels(M, N, DENSITY, panels=8, seed=0)
    return prob, info


def test_spmv_full_size_vs_scipy_and_adjoint(ctx, full):
    _, info = full
    a = info["A"]
    assert a.shape == (M, N) and abs(a.nnz - M * N * DENSITY) < 0.01 * M * N * DENSITY
    rng = np.random.default_rng(1)
    lm = LinearMap(ctx, LinearMapSpec.Sparse(a))
    x, z = rng.standard_normal(N), rng.standard_normal(N)
    y = rng.standard_normal(M)
    ax = lm.apply(x)
    ref = np.asarray(a @ x).ravel()
    assert np.abs(ax - ref).max() <= 1e-13 * np.abs(ref).max() * np.sqrt(a.nnz / M)
    t = lm.transpose()
    aty = t.apply(y)
    reft = np.asarray(a.T @ y).ravel()
    assert np.abs(aty - reft).max() <= 1e-13 * np.abs(reft).max() * np.sqrt(a.nnz / N)
    # <A x, y> = <x, A' y>
    lhs, rhs = float(ax @ y), float(x @ aty)
    assert lhs == pytest.approx(rhs, rel=1e-11)
    # A (2 x - 3 z) = 2 A x - 3 A z
    az = lm.apply(z)
    comb = lm.apply(2 * x - 3 * z)
    assert np.abs(comb - (2 * ax - 3 * az)).max() <= 1e-12 * np.abs(comb).max()
    # idempotence of the launch path: the same input gives the same bits (deterministic reductions)
    assert np.array_equal(lm.apply(x), ax)


def test_lasso_full_size_two_device_paths_agree(ctx, full, monkeypatch):
    prob, info = full
    with ProxADMMSolver(prob, {"cg_tol": 1e-12}, ctx=ctx) as s:
        x1 = s.solve()
        st1, tr1 = s.status.as_dict(), list(s.trace)
    monkeypatch.setenv("EPS_SPARSE_FORMAT", "csr")     # CSR kernels, CgStart/CgTail pair
    with ProxADMMSolver(prob, {"cg_tol": 1e-12}, ctx=ctx) as s:
        x2 = s.solve()
        st2, tr2 = s.status.as_dict(), list(s.trace)
    monkeypatch.delenv("EPS_SPARSE_FORMAT")
    assert st1["state"] == st2["state"] == 3                     # OPTIMAL
    assert st1["num_iterations"] == st2["num_iterations"]
    assert len(tr1) == len(tr2)
    for a, b in zip(tr1, tr2):
        for k in ("r_norm", "s_norm", "epsilon_primal", "epsilon_dual"):
            assert a[k] == pytest.approx(b[k], rel=1e-7), (k, a, b)
    xv1, xv2 = x1["x"], x2["x"]
    assert np.abs(xv1 - xv2).max() <= 1e-8 * max(1.0, np.abs(xv1).max())
    assert info["objective"](x1) == pytest.approx(info["objective"](x2), rel=1e-9)
    # optimality of the lasso at the ADMM tolerance: on the support the gradient of the smooth
    # part balances lam * sign(x); off the support it stays inside [-lam, lam]
    a, b, lam = info["A"], info["b"], info["lam"]
    g = 2 * np.asarray(a.T @ (np.asarray(a @ xv1).ravel() - b)).ravel()
    tol = 0.05 * lam                                              # rel_tol 1e-2 of the solver
    on = xv1 != 0
    assert on.sum() > 0
    assert np.abs(g[on] + lam * np.sign(xv1[on])).max() <= tol
    assert np.abs(g[~on]).max() <= lam + tol
