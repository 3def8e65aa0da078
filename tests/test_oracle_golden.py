"""Pins the CPU oracle (oracle/prox_admm_np.py, oracle/tv1d_dp.c) against outputs of the
unmodified reference solver committed under tests/golden/ (made by tests/golden/make_golden.py).

Runs on CPU.  The reference's own tests hold no golden vectors for this path, so these
fixtures -- the reference run on seeded inputs -- are what "the oracle follows the reference"
means here: same iteration counts, same per-epoch residuals, same solutions."""
import json
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")
sys.path.insert(0, GOLDEN)

import make_golden  # noqa: E402
from oracle import prox_admm_np as onp  # noqa: E402

INDEX = json.load(open(os.path.join(GOLDEN, "index.json")))
TRACE_KEYS = make_golden.TRACE_KEYS

# the biggest cases take the numpy oracle tens of seconds; keep the CPU suite to minutes
SLOW = {"lasso_config1_1000x5000"}


def check_trace(trace, gold_trace, rtol):
    assert len(trace) == gold_trace.shape[0]
    for t, g in zip(trace, gold_trace):
        gd = dict(zip(TRACE_KEYS, g))
        assert t["num_iterations"] == int(gd["num_iterations"])
        assert t["state"] == int(gd["state"])
        for k in TRACE_KEYS[2:]:
            assert t[k] == pytest.approx(gd[k], rel=rtol, abs=1e-13), (k, t, gd)


@pytest.mark.parametrize("name", sorted(INDEX["solve"]))
def test_generator_is_stable(name):
    """The seeded generators reproduce the exact bytes the reference consumed."""
    if name in SLOW:
        pytest.skip("covered by the GPU suite")
    prob, _, _ = make_golden.make_problem(name)
    assert make_golden.problem_digest(prob) == INDEX["solve"][name]["digest"]


@pytest.mark.parametrize("name", sorted(n for n in INDEX["solve"] if n not in SLOW))
def test_admm_matches_reference(name):
    prob, info, params = make_golden.make_problem(name)
    gold = np.load(os.path.join(GOLDEN, name + ".npz"))
    out = onp.ProxADMM(prob, params).solve()
    assert [out["m"], out["n"], out["N"]] == gold["mnN"].tolist()
    check_trace(out["trace"], gold["trace"], rtol=1e-9)
    st = dict(zip(TRACE_KEYS, gold["status"]))
    assert out["status"]["num_iterations"] == int(st["num_iterations"])
    assert out["status"]["state"] == int(st["state"])
    x = out["x"][INDEX["solve"][name]["solution_var"]]
    assert np.abs(x - gold["x"]).max() <= 1e-9 * max(1.0, np.abs(gold["x"]).max())
    assert info["objective"](out["x"]) == pytest.approx(float(gold["objective"]), rel=1e-9)


# bit-exact where the operator is compare/add/multiply only; tolerance where the reference
# sums or iterates (norm, Newton, Cholesky)
PROX_TOL = {"NORM_1": 0.0, "SUM_HINGE": 0.0, "SUM_DEADZONE": 0.0, "SUM_QUANTILE": 0.0,
            "NON_NEGATIVE": 0.0, "TOTAL_VARIATION_1D": 0.0, "NORM_2": 1e-15,
            "SUM_LOGISTIC": 1e-12, "SUM_EXP": 1e-10, "SUM_NEG_ENTR": 1e-10, "SUM_INV_POS": 1e-10,
            "SUM_NEG_LOG": 1e-15, "SUM_SQUARE": 1e-12}


@pytest.mark.parametrize("name", sorted(INDEX["prox"]))
def test_prox_matches_reference(name):
    f, lam, v = make_golden.prox_cases()[name]
    gold = np.load(os.path.join(GOLDEN, "prox_" + name + ".npz"))
    for k in v:
        assert np.array_equal(v[k], gold["v_" + k]), "generator drift"
    x = onp.eval_prox(f, lam, v)
    tol = 1e-10 if f.type in ("EXP", "LAMBDA_MAX", "NEG_LOG_DET", "NORM_NUCLEAR", "SEMIDEFINITE") else 1e-12 if (f.epigraph or f.type in ("MAX", "SUM_LARGEST", "SUM_KL_DIV", "SECOND_ORDER_CONE", "LOG_SUM_EXP")) else PROX_TOL[f.type]
    for k in v:
        g = gold["x_" + k]
        if tol == 0.0:
            assert np.array_equal(x[k], g), (name, np.abs(x[k] - g).max())
        else:
            assert np.abs(x[k] - g).max() <= tol * max(1.0, np.abs(g).max()), name


def test_tv1d_c_and_python_agree():
    rng = np.random.default_rng(5)
    for n in (1, 2, 3, 10, 257, 3000):
        y = np.repeat(rng.standard_normal(n // 50 + 1), 50)[:n] * 2 + rng.standard_normal(n)
        for lam in (0.0, 0.3, 5.0, 1e3):
            a = onp.tv1d(y, lam)
            b = onp.tv1d_py(y, lam)
            assert np.array_equal(a, b), (n, lam)


def test_tv1d_optimality():
    """KKT check of the dynamic program: y - x = lam * D'z with |z| <= 1, z = sign(Dx) off ties."""
    rng = np.random.default_rng(6)
    y = np.cumsum(rng.standard_normal(4000)) * 0.3
    lam = 2.0
    x = onp.tv1d(y, lam)
    z = np.cumsum(y - x)[:-1] / lam          # D'z = y - x  =>  z_k = sum_{i<=k}(y_i - x_i)/lam
    assert abs(np.sum(y - x)) < 1e-8
    assert np.all(np.abs(z) <= 1 + 1e-9)
    d = np.diff(x)
    nz = np.abs(d) > 1e-12
    assert np.allclose(z[nz], -np.sign(d[nz]), atol=1e-8) or np.allclose(z[nz], np.sign(d[nz]), atol=1e-8)


def test_scaled_zone_branches():
    v = np.array([-3.0, -1.0, -0.5, 0.0, 0.5, 1.0, 3.0])
    lam = np.full(v.size, 1.0)
    assert np.array_equal(onp.scaled_zone(v, lam, 1.0, 1.0, 0.0), [-2, 0, 0, 0, 0, 0, 2])      # soft threshold
    assert np.array_equal(onp.scaled_zone(v, lam, 1.0, 0.0, 0.0), [-3, -1, -0.5, 0, 0, 0, 2])  # hinge
    assert np.array_equal(onp.scaled_zone(v, lam, 1.0, 1.0, 0.5), [-2, -0.5, -0.5, 0, 0.5, 0.5, 2])


@pytest.mark.parametrize("name", sorted(INDEX.get("two_block", {})))
def test_two_block_matches_reference(name):
    """ProxADMMTwoBlockSolver restatement vs the stock reference solver (SolverParams.solver =
    PROX_ADMM_TWO_BLOCK); per-epoch residuals come from runs cut at max_iterations = k+1."""
    prob, info, params = make_golden.make_problem(name)
    assert make_golden.problem_digest(prob) == INDEX["two_block"][name]["digest"]
    gold = np.load(os.path.join(GOLDEN, name + ".npz"))
    out = onp.ProxADMMTwoBlock(prob, params).solve()
    check_trace(out["trace"], gold["trace"], rtol=1e-9)
    st = dict(zip(TRACE_KEYS, gold["status"]))
    assert out["status"]["num_iterations"] == int(st["num_iterations"])
    assert out["status"]["state"] == int(st["state"])
    x = out["x"][INDEX["two_block"][name]["solution_var"]]
    assert np.abs(x - gold["x"]).max() <= 1e-9 * max(1.0, np.abs(gold["x"]).max())
