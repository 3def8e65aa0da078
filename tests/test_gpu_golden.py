"""CUDA path (through the C ABI) against the committed outputs of the reference solver
(tests/golden/) and against the pinned CPU oracle on larger seeded inputs.

Bar (BASELINE.json north_star): identical iteration counts, residuals compared every epoch,
objective within 1e-6 relative in fp64; bit-exact for the compare/add-only prox operators."""
import json
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")
sys.path.insert(0, GOLDEN)

import make_golden  # noqa: E402
from epsilon_b200.problem import Affine, LinearMapSpec, ProxFunction, Term, Variable  # noqa: E402
from epsilon_b200.solver import ProxADMMSolver, eval_prox  # noqa: E402
from oracle import prox_admm_np as onp  # noqa: E402

pytestmark = pytest.mark.gpu

INDEX = json.load(open(os.path.join(GOLDEN, "index.json")))
TRACE_KEYS = make_golden.TRACE_KEYS

# kkt modes to exercise per case (the reference always factors; CG is the matrix-free mode
# used at full size and must reproduce the same stopping decisions)
MODES = {
    "lasso_dense_100x500": ["auto", "cg", "gram"],
    "lasso_dense_60x200": ["auto", "woodbury", "gram", "cg"],
    "lasso_sparse_4000x400": ["auto", "cg"],
    "lasso_sparse_30000x3000": ["auto", "cg"],
    "hinge_l2_sparse_5000x300": ["auto", "cg"],
    "logreg_l1_2000x200": ["auto", "cg"],
    "mv_lasso_40x100x3": ["auto", "cg", "gram"],
    "mv_lasso_120x60x4": ["auto", "cg"],
    "mv_lasso_sparse_3000x300x5": ["auto", "cg"],
    # nnz >= 2^18: packed as stream-SELL without being asked to
    "lasso_sparse_120000x4000": ["auto", "cg"],
    "logreg_l1_sparse_60000x2000": ["auto", "cg"],
    "hinge_l2_sparse_80000x3000": ["auto", "cg"],
}

# Sparse cases below the 2^18-nnz threshold are run a second time with EPS_SPARSE_FORMAT=sell, so that
# the kernels every headline number comes from (SellStreamKernel + SellReduceKernel, and in CG mode on
# the lasso the fused CgFusedTailKernel) are held to the REFERENCE's traces, not only to each other.
SELL_FORCED = ["lasso_sparse_4000x400", "lasso_sparse_30000x3000", "hinge_l2_sparse_5000x300",
               "logreg_l1_2000x200", "mv_lasso_sparse_3000x300x5"]
SELL_NATURAL = ["lasso_sparse_120000x4000", "logreg_l1_sparse_60000x2000", "hinge_l2_sparse_80000x3000"]


def _cases():
    for name in sorted(INDEX["solve"]):
        for mode in MODES.get(name, ["auto"]):
            yield name, mode, "default"
            if name in SELL_FORCED:
                yield name, mode, "sell"


_EXPECTED = {}


def _expected(name, prob, info, params):
    """The committed output of the reference for this case.  The fixtures hold outputs, not the
    (up to 40 MB) inputs; should the seeded generator produce different bytes on this host
    (numpy build differences), the pinned CPU oracle is run on the bytes at hand instead --
    the stop decision sits on a 0.3 % margin and only means something on identical inputs."""
    if name in _EXPECTED:
        return _EXPECTED[name]
    if make_golden.problem_digest(prob) == INDEX["solve"][name]["digest"]:
        gold = dict(np.load(os.path.join(GOLDEN, name + ".npz")))
    else:
        import warnings
        warnings.warn("generator drift for %s: checking against the CPU oracle" % name)
        o = onp.ProxADMM(prob, params).solve()
        gold = {"trace": np.array([[t[k] for k in TRACE_KEYS] for t in o["trace"]], dtype=np.float64),
                "status": np.array([o["status"][k] for k in TRACE_KEYS], dtype=np.float64),
                "mnN": np.array([o["m"], o["n"], o["N"]]),
                "x": o["x"][INDEX["solve"][name]["solution_var"]],
                "objective": np.float64(info["objective"](o["x"]))}
    _EXPECTED[name] = gold
    return gold


@pytest.mark.parametrize("name,mode,fmt", list(_cases()))
def test_admm_matches_reference_golden(ctx, name, mode, fmt, monkeypatch):
    prob, info, params = make_golden.make_problem(name)
    gold = _expected(name, prob, info, params)
    p = dict(params, kkt_mode=mode)
    if fmt != "default":
        monkeypatch.setenv("EPS_SPARSE_FORMAT", fmt)   # read by eps_map_csc at every ingest
    launches0 = ctx.kernel_launches
    with ProxADMMSolver(prob, p, ctx=ctx) as s:
        x = s.solve()
        st = s.status.as_dict()
        trace = list(s.trace)
        desc = s.describe()
    assert ctx.kernel_launches > launches0
    assert st["cg_unconverged"] == 0, desc     # every inner CG solve reached cg_tol (1e-12)
    if (fmt == "sell" or name in SELL_NATURAL) and mode == "cg" and name.startswith("lasso_sparse"):
        # the headline kernel trio: stream-SELL passes + the fused reduce / CG-update kernel
        assert "fused-matvec" in desc and "fused-tail" in desc, desc
    assert [st["m"], st["n"], st["N"]] == gold["mnN"].tolist(), desc
    assert len(trace) == gold["trace"].shape[0], (len(trace), gold["trace"].shape, desc)
    rtol = 1e-9 if mode != "cg" else 1e-7
    for t, g in zip(trace, gold["trace"]):
        gd = dict(zip(TRACE_KEYS, g))
        assert t["num_iterations"] == int(gd["num_iterations"])
        assert t["state"] == int(gd["state"]), (t, gd, desc)
        for k in TRACE_KEYS[2:]:
            assert t[k] == pytest.approx(gd[k], rel=rtol, abs=1e-12), (k, t, gd, desc)
    gst = dict(zip(TRACE_KEYS, gold["status"]))
    assert st["num_iterations"] == int(gst["num_iterations"])
    assert st["state"] == int(gst["state"])
    v = INDEX["solve"][name]["solution_var"]
    assert np.abs(x[v] - gold["x"]).max() <= 1e-7 * max(1.0, np.abs(gold["x"]).max())
    assert info["objective"](x) == pytest.approx(float(gold["objective"]), rel=1e-6)


PROX_TOL = {"NORM_1": 0.0, "SUM_HINGE": 0.0, "SUM_DEADZONE": 0.0, "SUM_QUANTILE": 0.0,
            "NON_NEGATIVE": 0.0, "TOTAL_VARIATION_1D": 1e-13, "NORM_2": 1e-14,
            "SUM_LOGISTIC": 1e-12, "SUM_EXP": 1e-10, "SUM_NEG_ENTR": 1e-10, "SUM_INV_POS": 1e-10,
            "SUM_NEG_LOG": 1e-15, "SUM_SQUARE": 1e-10}


@pytest.mark.parametrize("name", sorted(INDEX["prox"]))
def test_prox_matches_reference_golden(ctx, name):
    f, lam, v = make_golden.prox_cases()[name]
    gold = np.load(os.path.join(GOLDEN, "prox_" + name + ".npz"))
    x = eval_prox(f, lam, v, ctx=ctx)
    tol = 1e-10 if f.type in ("EXP", "LAMBDA_MAX", "NEG_LOG_DET", "NORM_NUCLEAR", "SEMIDEFINITE") else 1e-12 if (f.epigraph or f.type in ("MAX", "SUM_LARGEST", "SUM_KL_DIV", "SECOND_ORDER_CONE", "LOG_SUM_EXP")) else PROX_TOL[f.type]
    for k in v:
        g = gold["x_" + k]
        if tol == 0.0:
            assert np.array_equal(x[k], g), (name, np.abs(x[k] - g).max())
        else:
            assert np.abs(x[k] - g).max() <= tol * max(1.0, np.abs(g).max()), (name, np.abs(x[k] - g).max())


def _simple(ptype, n, alpha=1.0, **kw):
    return ProxFunction(ptype, args=[Affine(n, [Term(Variable("x", n))])], alpha=alpha, **kw)


@pytest.mark.parametrize("ptype,kw", [
    ("NORM_1", {}), ("SUM_HINGE", {}), ("SUM_DEADZONE", {"zone_m": 0.25}), ("NON_NEGATIVE", {}),
])
@pytest.mark.parametrize("n", [1, 31, 1 << 20, (1 << 22) + 3])
def test_elementwise_prox_bit_exact_vs_oracle(ctx, ptype, kw, n):
    """Ragged and large sizes: 0 ulp against the pinned oracle (compare/add/mul only)."""
    rng = np.random.default_rng(n)
    v = rng.standard_normal(n) * 3
    f = _simple(ptype, n, alpha=0.9, **kw)
    x = eval_prox(f, 1.7, {"x": v}, ctx=ctx)["x"]
    ref = onp.eval_prox(f, 1.7, {"x": v})["x"]
    assert np.array_equal(x, ref)


def test_quantile_vector_bit_exact(ctx):
    n = 100003
    rng = np.random.default_rng(2)
    f = _simple("SUM_QUANTILE", n, quantile_alpha=rng.random(n), quantile_beta=rng.random(n))
    v = rng.standard_normal(n) * 2
    assert np.array_equal(eval_prox(f, 0.8, {"x": v}, ctx=ctx)["x"], onp.eval_prox(f, 0.8, {"x": v})["x"])


@pytest.mark.parametrize("n", [7, 4096, 300001])
def test_logistic_newton_vs_oracle(ctx, n):
    """Per-element Newton on the device vs the reference's global-line-search Newton:
    both solve x - v + lam*sigma(x) = 0; agreement to 1e-12 absolute."""
    rng = np.random.default_rng(n)
    v = rng.standard_normal(n) * 5
    f = _simple("SUM_LOGISTIC", n)
    x = eval_prox(f, 2.5, {"x": v}, ctx=ctx)["x"]
    ref = onp.eval_prox(f, 2.5, {"x": v})["x"]
    assert np.abs(x - ref).max() <= 1e-12
    # and the optimality condition itself
    resid = x - v + 2.5 / (1 + np.exp(-x))
    assert np.abs(resid).max() <= 1e-12


def test_logistic_extreme_inputs_are_finite(ctx):
    """The reference's exp() overflows for |x| > 709 (sum_logistic.cc:14,22,29); the device
    kernel uses a stable sigmoid and must return the exact limit values."""
    v = np.array([-1e4, -800.0, -50.0, 0.0, 50.0, 800.0, 1e4])
    x = eval_prox(_simple("SUM_LOGISTIC", v.size), 2.0, {"x": v}, ctx=ctx)["x"]
    assert np.all(np.isfinite(x))
    # sigma underflows to 0: x = v up to the rounding of EvalProx's 1/sqrt(lambda) scaling
    assert x[0] == pytest.approx(-1e4, rel=1e-15) and x[1] == pytest.approx(-800.0, rel=1e-15)
    assert x[-1] == pytest.approx(1e4 - 2.0, abs=1e-9) and x[-2] == pytest.approx(798.0, abs=1e-9)


@pytest.mark.parametrize("lam", [1e-3, 0.5, 7.0, 300.0, 1e5, 3e8, 1e14, 1e20])
def test_logistic_prox_wide_range_vs_high_precision_root(ctx, lam):
    """The three-stage SUM_LOGISTIC kernel (fp32 monotone Newton, one fp64 exp + Taylor-carried steps, damped
    Newton of newton.cc:47-91 as the fallback) against the root of x - v + lam*sigma(x) found independently by
    bisection in extended precision.  Large lam makes part of every warp take the fallback."""
    rng = np.random.default_rng(int(lam * 1000) % 9973)
    v = np.concatenate([rng.standard_normal(6000) * 3, rng.uniform(-80, 80, 3000), rng.standard_normal(1000) * 1e3,
                        np.array([0.0, 0.5 * lam, -0.5 * lam, lam, 1e-300, -1e-300])])
    x = eval_prox(_simple("SUM_LOGISTIC", v.size), lam, {"x": v}, ctx=ctx)["x"]
    vl, ll = v.astype(np.longdouble), np.longdouble(lam)
    lo, hi = vl - ll, vl.copy()
    for _ in range(160):
        mid = (lo + hi) / 2
        e = np.exp(-np.abs(mid))
        s = np.where(mid >= 0, 1 / (1 + e), e / (1 + e))
        pos = mid - vl + ll * s > 0
        hi = np.where(pos, mid, hi)
        lo = np.where(pos, lo, mid)
    root = ((lo + hi) / 2).astype(np.float64)
    assert np.all(np.isfinite(x))
    # conditioning: the residual x - v + lam*sigma(x) cancels terms of size |v| + lam*sigma in fp64 (for the
    # reference as well), which moves the computed root by that rounding over f' = 1 + lam*sigma'
    e = np.exp(-np.abs(root))
    sg = np.where(root >= 0, 1 / (1 + e), e / (1 + e))
    cond = 8 * np.finfo(np.float64).eps * (np.abs(v) + lam * sg) / (1 + lam * sg * (1 - sg))
    assert np.all(np.abs(x - root) <= 5e-14 * np.maximum(1.0, np.abs(root)) + cond)


@pytest.mark.parametrize("lam", [1e-3, 0.4, 40.0, 1e4])
@pytest.mark.parametrize("n", [2047, 2048, 2049, 4096, 4097, 6145, 100003])
def test_tv1d_tile_edges_and_group_regimes_vs_oracle(ctx, n, lam):
    """The tiled active-set solver (tv1d_par.cu: tiles of 2048 boundaries, carries between tiles, dirty-tile lists)
    at lengths around the tile size, from 'every boundary active' (tiny lambda) to 'one group across all tiles'
    (huge lambda), against the dynamic program (oracle/tv1d_dp.c = glmgen tf_dp)."""
    rng = np.random.default_rng(n + int(lam * 1000))
    v = np.repeat(rng.standard_normal(n // 300 + 1) * 2, 300)[:n] + 0.5 * rng.standard_normal(n)
    f = _simple("TOTAL_VARIATION_1D", n, alpha=1.0)
    x = eval_prox(f, lam, {"x": v}, ctx=ctx)["x"]
    ref = onp.eval_prox(f, lam, {"x": v})["x"]
    assert np.abs(x - ref).max() <= 1e-11 * max(1.0, np.abs(ref).max())


@pytest.mark.parametrize("n", [1, 2, 3, 1000, 65537, 1 << 20])
def test_tv1d_vs_oracle(ctx, n):
    rng = np.random.default_rng(n)
    v = np.repeat(rng.standard_normal(n // 200 + 1) * 3, 200)[:n] + rng.standard_normal(n)
    f = _simple("TOTAL_VARIATION_1D", n, alpha=3.0)
    x = eval_prox(f, 1.5, {"x": v}, ctx=ctx)["x"]
    ref = onp.eval_prox(f, 1.5, {"x": v})["x"]
    assert np.abs(x - ref).max() <= 1e-11 * max(1.0, np.abs(ref).max())


def test_norm2_groups_vs_oracle(ctx):
    rng = np.random.default_rng(3)
    for n, alpha in [(1, 0.5), (50, 2.0), (50, 1e3), (5000, 3.0), (70000, 10.0)]:
        v = rng.standard_normal(n)
        f = _simple("NORM_2", n, alpha=alpha)
        x = eval_prox(f, 1.0, {"x": v}, ctx=ctx)["x"]
        ref = onp.eval_prox(f, 1.0, {"x": v})["x"]
        assert np.abs(x - ref).max() <= 1e-14 * max(1.0, np.abs(ref).max())


def test_unsupported_structure_is_an_error(ctx):
    """A reference CHECK (vector_prox.cc:118-120: argument not scalar/diagonal) surfaces as
    an error code + message, never as a wrong answer or a fallback."""
    from epsilon_b200.capi import EpsError
    rng = np.random.default_rng(0)
    x = Variable("x", 5)
    f = ProxFunction("NORM_1", args=[Affine(4, [Term(x, LinearMapSpec.Dense(rng.standard_normal((4, 5))))])])
    with pytest.raises(EpsError):
        eval_prox(f, 1.0, {"x": rng.standard_normal(5)}, ctx=ctx)


def _tb_cases():
    for name in sorted(INDEX.get("two_block", {})):
        yield name, "default"
        if name in ("tb_lasso_sparse_4000x400", "tb_logreg_l1_2000x200"):
            yield name, "sell"


@pytest.mark.parametrize("name,fmt", list(_tb_cases()))
def test_two_block_matches_reference_golden(ctx, name, fmt, monkeypatch):
    """ProxADMMTwoBlockSolver (SolverParams.solver = PROX_ADMM_TWO_BLOCK) on the device vs the
    stock reference solver: identical iteration counts, per-epoch residuals, solution."""
    prob, info, params = make_golden.make_problem(name)
    if make_golden.problem_digest(prob) == INDEX["two_block"][name]["digest"]:
        gold = dict(np.load(os.path.join(GOLDEN, name + ".npz")))
    else:
        o = onp.ProxADMMTwoBlock(prob, params).solve()
        gold = {"trace": np.array([[t[k] for k in TRACE_KEYS] for t in o["trace"]], dtype=np.float64),
                "status": np.array([o["status"][k] for k in TRACE_KEYS], dtype=np.float64),
                "x": o["x"][INDEX["two_block"][name]["solution_var"]],
                "objective": np.float64(info["objective"](o["x"]))}
    if fmt != "default":
        monkeypatch.setenv("EPS_SPARSE_FORMAT", fmt)
    with ProxADMMSolver(prob, dict(params, solver="PROX_ADMM_TWO_BLOCK"), ctx=ctx) as s:
        x = s.solve()
        st = s.status.as_dict()
        trace = list(s.trace)
        desc = s.describe()
    assert len(trace) == gold["trace"].shape[0], (len(trace), gold["trace"].shape, desc)
    for t, g in zip(trace, gold["trace"]):
        gd = dict(zip(TRACE_KEYS, g))
        assert t["num_iterations"] == int(gd["num_iterations"]) and t["state"] == int(gd["state"]), (t, gd, desc)
        for k in TRACE_KEYS[2:]:
            assert t[k] == pytest.approx(gd[k], rel=1e-8, abs=1e-12), (k, t, gd, desc)
    gst = dict(zip(TRACE_KEYS, gold["status"]))
    assert st["num_iterations"] == int(gst["num_iterations"]) and st["state"] == int(gst["state"])
    v = INDEX["two_block"][name]["solution_var"]
    assert np.abs(x[v] - gold["x"]).max() <= 1e-7 * max(1.0, np.abs(gold["x"]).max())
    assert info["objective"](x) == pytest.approx(float(gold["objective"]), rel=1e-6)


def _mat(ptype, m, n, epigraph=False, alpha=1.0):
    args = [Affine(m * n, [Term(Variable("x", m * n))])]
    if epigraph:
        args.append(Affine(1, [Term(Variable("t", 1))]))
    return ProxFunction(ptype, args=args, alpha=alpha, epigraph=epigraph, mat_dims=(m, n))


@pytest.mark.parametrize("ptype,m,n,epigraph", [
    ("SEMIDEFINITE", 1, 1, False), ("SEMIDEFINITE", 2, 2, False), ("SEMIDEFINITE", 33, 33, False),
    ("SEMIDEFINITE", 150, 150, False),      # above 112 the Jacobi sweeps run out of global memory
    ("NEG_LOG_DET", 127, 127, False), ("LAMBDA_MAX", 45, 45, False), ("LAMBDA_MAX", 131, 131, True),
    ("NORM_NUCLEAR", 200, 130, False), ("NORM_NUCLEAR", 60, 60, True), ("NORM_NUCLEAR", 7, 1, False),
])
def test_matrix_functions_vs_oracle(ctx, ptype, m, n, epigraph):
    """Jacobi eigen-decomposition on the device against LAPACK in the oracle: sizes off the fixtures
    (1 x 1, odd n, the shared- and the global-memory variant of the sweep)."""
    rng = np.random.default_rng(m * 1000 + n)
    y = rng.standard_normal(m * n)
    if ptype == "NEG_LOG_DET":
        y = (y.reshape(m, n) + 3 * np.eye(n)).ravel()
    v = {"x": y}
    if epigraph:
        v["t"] = np.array([1.5])
    f = _mat(ptype, m, n, epigraph, alpha=0.8)
    x = eval_prox(f, 1.3, v, ctx=ctx)
    ref = onp.eval_prox(f, 1.3, v)
    for k in v:
        assert np.abs(x[k] - ref[k]).max() <= 1e-10 * max(1.0, np.abs(ref[k]).max()), (k, np.abs(x[k] - ref[k]).max())


def test_semidefinite_projection_properties(ctx):
    """Projection onto the PSD cone: symmetric part is PSD, the skew part is kept, idempotent; a
    rank-deficient input (repeated / zero eigenvalues) converges."""
    n = 96
    rng = np.random.default_rng(77)
    b = rng.standard_normal((n, 5))
    y = b @ b.T - 0.5 * np.eye(n) + np.triu(rng.standard_normal((n, n)), 1) * 0.1
    f = _mat("SEMIDEFINITE", n, n)
    x = eval_prox(f, 1.0, {"x": y.ravel(order="F")}, ctx=ctx)["x"].reshape((n, n), order="F")
    sym, skew = (x + x.T) / 2, (x - x.T) / 2
    assert np.linalg.eigvalsh(sym).min() >= -1e-12
    assert np.abs(skew - (y - y.T) / 2).max() <= 1e-14
    x2 = eval_prox(f, 1.0, {"x": x.ravel(order="F")}, ctx=ctx)["x"].reshape((n, n), order="F")
    assert np.abs(x2 - x).max() <= 1e-11


def test_nuclear_norm_wide_argument_is_rejected(ctx):
    """m < n: the reference's nested operator is sized min(m, n) for n singular values
    (ortho_invariant.cc:77 vs :41-46); refused here instead of imitated."""
    from epsilon_b200.capi import EpsError
    with pytest.raises(EpsError):
        eval_prox(_mat("NORM_NUCLEAR", 6, 14), 1.0, {"x": np.ones(84)}, ctx=ctx)
