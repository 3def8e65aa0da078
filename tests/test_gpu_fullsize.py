"""Parity at BASELINE.json's full size (sparse lasso 1M x 100k, nnz 1e8), where the compiled
reference cannot run (it would form and invert a dense 100k x 100k matrix,
src/epsilon/linear/sparse_matrix_impl.cc:60-78): the linear maps are checked against scipy on
the same bytes and through size-independent properties (adjointness, linearity); the whole
solve -- stream-SELL kernels + fused CG update, and plain CSR kernels + unfused CG -- against
tests/golden/lasso_sparse_1Mx100k.npz: the pinned CPU oracle run matrix-free on exactly this
instance (tests/golden/make_fullsize.py), per-epoch residuals, iteration count, solution and
objective; and the optimality conditions of the lasso."""
import hashlib
import json
import os

import numpy as np
import pytest

from epsilon_b200 import problems
from epsilon_b200.problem import LinearMapSpec
from epsilon_b200.solver import LinearMap, ProxADMMSolver

pytestmark = pytest.mark.gpu

M, N, DENSITY = 1_000_000, 100_000, 1e-3
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TRACE_KEYS = ("state", "num_iterations", "r_norm", "s_norm", "epsilon_primal", "epsilon_dual")


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


def _check_against_fixture(x, st, trace, info, desc):
    gold = np.load(os.path.join(GOLDEN, "lasso_sparse_1Mx100k.npz"))
    assert [st["m"], st["n"], st["N"]] == gold["mnN"].tolist(), desc
    assert len(trace) == gold["trace"].shape[0], (len(trace), desc)
    for t, g in zip(trace, gold["trace"]):
        gd = dict(zip(TRACE_KEYS, g))
        assert t["num_iterations"] == int(gd["num_iterations"]) and t["state"] == int(gd["state"]), (t, gd)
        for k in TRACE_KEYS[2:]:
            assert t[k] == pytest.approx(gd[k], rel=1e-7), (k, t, gd, desc)
    gst = dict(zip(TRACE_KEYS, gold["status"]))
    assert st["num_iterations"] == int(gst["num_iterations"]) and st["state"] == int(gst["state"]) == 3
    assert st["cg_unconverged"] == 0
    xv = x["x"]
    assert np.abs(xv - gold["x"]).max() <= 1e-7 * max(1.0, np.abs(gold["x"]).max())
    assert info["objective"](x) == pytest.approx(float(gold["objective"]), rel=1e-9)


def test_generator_reproduces_the_fixture_inputs(full):
    """The oracle fixture only means something on identical input bytes."""
    _, info = full
    meta = json.load(open(os.path.join(GOLDEN, "lasso_sparse_1Mx100k.json")))
    a = info["A"]
    h = hashlib.sha256()
    h.update(a.indptr.astype(np.int32).tobytes())
    h.update(a.indices.astype(np.int32).tobytes())
    h.update(a.data.astype(np.float64).tobytes())
    h.update(np.asarray(info["b"], dtype=np.float64).tobytes())
    h.update(np.float64(info["lam"]).tobytes())
    assert h.hexdigest() == meta["digest"]


@pytest.mark.parametrize("fmt", ["sell", "csr"])
def test_lasso_full_size_matches_oracle_fixture(ctx, full, fmt, monkeypatch):
    """sell: SellStreamKernel + SellReduceKernel + CgFusedTailKernel (the kernels bench.py times);
    csr: CsrSpmvKernel + CgStartKernel/CgTailKernel."""
    prob, info = full
    monkeypatch.setenv("EPS_SPARSE_FORMAT", fmt)
    with ProxADMMSolver(prob, {"cg_tol": 1e-12}, ctx=ctx) as s:
        x = s.solve()
        st, trace, desc = s.status.as_dict(), list(s.trace), s.describe()
    assert ("fused-tail" in desc) == (fmt == "sell"), desc
    _check_against_fixture(x, st, trace, info, desc)
    if fmt == "csr":
        return
    # optimality of the lasso at the ADMM tolerance: on the support the gradient of the smooth
    # part balances lam * sign(x); off the support it stays inside [-lam, lam]
    xv = x["x"]
    a, b, lam = info["A"], info["b"], info["lam"]
    g = 2 * np.asarray(a.T @ (np.asarray(a @ xv).ravel() - b)).ravel()
    tol = 0.05 * lam                                              # rel_tol 1e-2 of the solver
    on = xv != 0
    assert on.sum() > 0
    assert np.abs(g[on] + lam * np.sign(xv[on])).max() <= tol
    assert np.abs(g[~on]).max() <= lam + tol
