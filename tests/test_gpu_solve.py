"""End-to-end parity of the device ProxADMM loop with the compiled reference solver on the
same input bytes: identical iteration counts, per-epoch residuals, objective within 1e-6."""
import numpy as np
import pytest

from epsilon_b200 import problems
from epsilon_b200.solver import ProxADMMSolver

pytestmark = pytest.mark.gpu

RES_KEYS = ("r_norm", "s_norm", "epsilon_primal", "epsilon_dual")


def compare(ctx, ref, prob, info, params=None, res_rtol=1e-9, obj_rtol=1e-6, max_iterations=None):
    ref_params = {}
    if max_iterations:
        ref_params["max_iterations"] = max_iterations
    r = ref.run(prob, params=ref_params, mode="trace")
    p = dict(params or {})
    if max_iterations:
        p["max_iterations"] = max_iterations
    with ProxADMMSolver(prob, p, ctx=ctx) as s:
        x = s.solve()
        st = s.status.as_dict()
        trace = list(s.trace)
        desc = s.describe()
    assert st["m"] == int(r["m"]) and st["n"] == int(r["n"]) and st["N"] == int(r["N"]), desc
    assert len(trace) == len(r["trace"]), (len(trace), len(r["trace"]), desc)
    for a, b in zip(trace, r["trace"]):
        assert a["num_iterations"] == b["num_iterations"]
        for k in RES_KEYS:
            assert a[k] == pytest.approx(b[k], rel=res_rtol, abs=1e-13), (k, a, b, desc)
        assert a["state"] == b["state"]
    assert st["state"] == r["status"]["state"]
    assert st["num_iterations"] == r["status"]["num_iterations"]
    obj, obj_ref = info["objective"](x), info["objective"](r["x"])
    assert obj == pytest.approx(obj_ref, rel=obj_rtol)
    v = info["solution_var"]
    assert np.abs(x[v] - r["x"][v]).max() <= 1e-7 * max(1.0, np.abs(r["x"][v]).max())
    return st


def test_lasso_dense_woodbury(ctx, ref):
    prob, info = problems.lasso(100, 500)
    compare(ctx, ref, prob, info)


@pytest.mark.parametrize("mode", ["gram", "cg", "woodbury"])
def test_lasso_dense_modes(ctx, ref, mode):
    prob, info = problems.lasso(60, 200, seed=1)
    compare(ctx, ref, prob, info, params={"kkt_mode": mode}, res_rtol=1e-8)


def test_lasso_dense_tall_gram(ctx, ref):
    prob, info = problems.lasso(600, 150, seed=2)
    compare(ctx, ref, prob, info)


def test_lasso_config1_full(ctx, ref):
    """BASELINE config 1: dense lasso m=1000 n=5000 (reference: 40 iterations)."""
    prob, info = problems.lasso(1000, 5000)
    st = compare(ctx, ref, prob, info)
    assert st["state"] == 3


@pytest.mark.parametrize("mode", ["gram", "cg"])
def test_lasso_sparse(ctx, ref, mode):
    prob, info = problems.lasso(4000, 400, density=0.02, seed=3)
    compare(ctx, ref, prob, info, params={"kkt_mode": mode}, res_rtol=1e-8)


def test_logreg_l1(ctx, ref):
    prob, info = problems.logreg_l1(2000, 200, density=0.05, seed=4)
    compare(ctx, ref, prob, info, res_rtol=1e-7, max_iterations=200)


def test_hinge_l2(ctx, ref):
    prob, info = problems.hinge_l2(1500, 100, seed=5)
    compare(ctx, ref, prob, info, res_rtol=1e-8, max_iterations=300)


def test_group_lasso(ctx, ref):
    prob, info = problems.group_lasso(400, 10, 12, seed=6)
    compare(ctx, ref, prob, info, res_rtol=1e-8, max_iterations=300)


def test_tv_1d(ctx, ref):
    prob, info = problems.tv_1d(5000, seed=7)
    compare(ctx, ref, prob, info, res_rtol=1e-8, max_iterations=300)


def test_set_constant_on_an_operator_built_without_one(ctx):
    """eps_solver_set_constant (Solver::SetParameterValue, solver.cc:109-116) gives an elementwise operator
    that was built with g = 0 a constant later: NORM_1(x) -> NORM_1(x + c).  Must equal a solver built
    with c from the start, and the pinned CPU oracle on that problem."""
    from epsilon_b200.problem import Term
    from oracle import prox_admm_np as onp
    rng = np.random.default_rng(31)
    prob, info = problems.lasso(300, 120, seed=9)
    c = 0.05 * rng.standard_normal(120)
    with ProxADMMSolver(prob, {}, ctx=ctx) as s:
        s.solve()
        s.set_constant(1, 0, c)              # argument 0 of prox function 1 (NORM_1)
        s.set_constant(0, 0, -info["b"])     # unchanged value of the SUM_SQUARE constant: c0 recomputed
        x1 = s.solve()                       # warm_start = 0: a cold solve of the new problem
        st1, tr1 = s.status.as_dict(), list(s.trace)
    prob2, _ = problems.lasso(300, 120, seed=9)
    prob2.prox[1].args[0].terms.append(Term(c, name="c"))
    with ProxADMMSolver(prob2, {}, ctx=ctx) as s:
        x2 = s.solve()
        st2, tr2 = s.status.as_dict(), list(s.trace)
    assert st1["num_iterations"] == st2["num_iterations"] and len(tr1) == len(tr2)
    for a, b in zip(tr1, tr2):
        for k in RES_KEYS:
            assert a[k] == b[k]
    assert np.array_equal(x1["x"], x2["x"])
    o = onp.ProxADMM(prob2).solve()
    assert o["status"]["num_iterations"] == st2["num_iterations"]
    assert np.abs(o["x"]["x"] - x2["x"]).max() <= 1e-9
