#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the UNMODIFIED reference solver.

Run in the build container (needs oracle/_ref/epsilon_ref, i.e. /root/reference + `make -C
oracle`):   python tests/golden/make_golden.py

Each fixture records, for one seeded problem of epsilon_b200.problems (the same bytes go to
the reference and, in the tests, to the numpy oracle and the CUDA path):
  * a SHA-256 over the generated input buffers (guards against generator drift),
  * the reference's per-epoch residual trace, final status, m/n/N and the solution x,
and for single operators (the harness' `prox` mode = EvalProx, solvemodule.cc:189-242) the
input v, lambda and the reference's output.
The reference holds no golden vectors of its own for this path; these
outputs of the reference itself are what pins the oracle.
"""
from __future__ import annotations

import hashlib
import json
import os
import sys

import numpy as np
import scipy.sparse as sp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from epsilon_b200 import problems  # noqa: E402
from epsilon_b200.problem import Affine, LinearMapSpec, ProxFunction, Term, Variable  # noqa: E402

# name -> (generator, args, kwargs, solver params)
SOLVE_CASES = {
    "lasso_dense_100x500": ("lasso", (100, 500), {}, {}),
    "lasso_dense_tall_600x150": ("lasso", (600, 150), {"seed": 2}, {}),
    "lasso_dense_60x200": ("lasso", (60, 200), {"seed": 1}, {}),
    "lasso_config1_1000x5000": ("lasso", (1000, 5000), {}, {}),
    "lasso_sparse_4000x400": ("lasso", (4000, 400), {"density": 0.02, "seed": 3}, {}),
    "lasso_sparse_30000x3000": ("lasso", (30000, 3000), {"density": 0.001, "seed": 8}, {}),
    # at / above the 2^18-nnz threshold from which the device packs sparse maps as stream-SELL (the
    # headline kernels), still small enough for the reference to factor (n^2 dense inverse)
    "lasso_sparse_120000x4000": ("lasso", (120000, 4000), {"density": 0.001, "seed": 13}, {}),
    "logreg_l1_sparse_60000x2000": ("logreg_l1", (60000, 2000), {"density": 0.005, "seed": 14}, {"max_iterations": 200}),
    "hinge_l2_sparse_80000x3000": ("hinge_l2", (80000, 3000), {"density": 0.002, "seed": 15}, {"max_iterations": 300}),
    "logreg_l1_2000x200": ("logreg_l1", (2000, 200), {"density": 0.05, "seed": 4}, {"max_iterations": 200}),
    "logreg_l1_dense_500x80": ("logreg_l1", (500, 80), {"seed": 9}, {"max_iterations": 200}),
    "hinge_l2_1500x100": ("hinge_l2", (1500, 100), {"seed": 5}, {"max_iterations": 300}),
    "hinge_l2_sparse_5000x300": ("hinge_l2", (5000, 300), {"density": 0.02, "seed": 10}, {"max_iterations": 300}),
    "group_lasso_400x10x12": ("group_lasso", (400, 10, 12), {"seed": 6}, {"max_iterations": 300}),
    "tv_1d_5000": ("tv_1d", (5000,), {"seed": 7}, {"max_iterations": 300}),
    "tv_1d_200000": ("tv_1d", (200000,), {"seed": 11}, {"max_iterations": 100}),
    # stop by max_iterations (not a multiple of the epoch): MAX_ITERATIONS_REACHED path
    "lasso_maxiter_17": ("lasso", (80, 120), {"seed": 12}, {"max_iterations": 17}),
    # epigraph form (use_epigraph=true): NORM_1 epigraph operator over (z, t) + AFFINE on a copy of t
    # matrix variables: A*X = (I_k (x) A) vec(X), KroneckerProductImpl (benchmark.py:46 "mv_lasso")
    "mv_lasso_40x100x3": ("mv_lasso", (40, 100, 3), {"seed": 2}, {}),
    "mv_lasso_120x60x4": ("mv_lasso", (120, 60, 4), {"seed": 2}, {}),
    "mv_lasso_sparse_3000x300x5": ("mv_lasso", (3000, 300, 5), {"density": 0.02, "seed": 3}, {}),
    "l1_epigraph_denoise_300": ("l1_epigraph_denoise", (300,), {"seed": 1}, {"max_iterations": 400}),
    # VectorProx::InitDiagonal (weights with exact zeros) and has_axis (row norms of a matrix variable)
    "weighted_lasso_200x300": ("weighted_lasso", (200, 300), {"seed": 16}, {"max_iterations": 300}),
    "weighted_lasso_sparse_4000x400": ("weighted_lasso", (4000, 400), {"density": 0.02, "seed": 17}, {"max_iterations": 300}),
    "mv_row_group_lasso_60x80x4": ("mv_row_group_lasso", (60, 80, 4), {"seed": 18}, {"max_iterations": 300}),
    "nuclear_denoise_30x12": ("nuclear_denoise", (30, 12), {"seed": 2}, {"max_iterations": 400, "rel_tol": 1e-5, "abs_tol": 1e-6}),
    "covsel_12": ("sparse_inverse_covariance", (12,), {"seed": 3}, {"max_iterations": 600, "rel_tol": 1e-5, "abs_tol": 1e-6}),
}


# ProxADMMTwoBlockSolver (SolverParams.solver = PROX_ADMM_TWO_BLOCK, solve_test.py:59)
TWO_BLOCK_CASES = {
    "tb_lasso_dense_100x500": ("lasso", (100, 500), {}, {"max_iterations": 500}),
    "tb_lasso_sparse_4000x400": ("lasso", (4000, 400), {"density": 0.02, "seed": 3}, {"max_iterations": 500}),
    "tb_logreg_l1_2000x200": ("logreg_l1", (2000, 200), {"density": 0.05, "seed": 4}, {"max_iterations": 300}),
    "tb_hinge_l2_1500x100": ("hinge_l2", (1500, 100), {"seed": 5}, {"max_iterations": 300}),
    "tb_group_lasso_400x10x12": ("group_lasso", (400, 10, 12), {"seed": 6}, {"max_iterations": 300}),
    "tb_tv_1d_5000": ("tv_1d", (5000,), {"seed": 7}, {"max_iterations": 300}),
    "tb_lasso_maxiter_13": ("lasso", (80, 120), {"seed": 12}, {"max_iterations": 13}),
}


def problem_digest(prob) -> str:
    """SHA-256 over every constant and matrix buffer of the problem, in traversal order."""
    h = hashlib.sha256()

    def feed_map(m):
        if m is None:
            return
        if m.dense is not None:
            h.update(np.asfortranarray(m.dense).tobytes(order="F"))
        if m.sparse is not None:
            a = sp.csc_matrix(m.sparse)
            h.update(a.indptr.astype(np.int32).tobytes())
            h.update(a.indices.astype(np.int32).tobytes())
            h.update(a.data.astype(np.float64).tobytes())
        if m.diag is not None:
            h.update(np.asarray(m.diag, dtype=np.float64).tobytes())
        h.update(np.float64(m.scalar).tobytes())
        for k in (m.kron or ()):
            feed_map(k)

    for aff in [a for f in prob.prox for a in f.args] + list(prob.constraints):
        for t in aff.terms:
            feed_map(t.map)
            if not isinstance(t.operand, Variable):
                h.update(np.asarray(t.operand, dtype=np.float64).tobytes())
    for f in prob.prox:
        h.update(np.float64(f.alpha).tobytes())
    return h.hexdigest()


def make_problem(name):
    gen, args, kw, params = (TWO_BLOCK_CASES if name.startswith("tb_") else SOLVE_CASES)[name]
    prob, info = getattr(problems, gen)(*args, **kw)
    return prob, info, params


TRACE_KEYS = ("state", "num_iterations", "r_norm", "s_norm", "epsilon_primal", "epsilon_dual")


# ---- single operators ------------------------------------------------------------------
def prox_cases():
    """name -> (ProxFunction, lambda, {var: v})."""
    rng = np.random.default_rng(1234)
    out = {}

    def simple(ptype, n, alpha=1.0, scale=1.0, **kw):
        x = Variable("x", n)
        args = [Affine(n, [Term(x) if scale == 1.0 else Term(x, LinearMapSpec.Scalar(scale, n))])]
        return ProxFunction(ptype, args=args, alpha=alpha, **kw)

    n = 2000
    out["norm_1"] = (simple("NORM_1", n, alpha=0.7), 1.3, {"x": rng.standard_normal(n) * 2})
    out["norm_1_scaled_arg"] = (simple("NORM_1", n, alpha=0.4, scale=-2.5), 0.6, {"x": rng.standard_normal(n)})
    out["sum_hinge"] = (simple("SUM_HINGE", n), 0.8, {"x": rng.standard_normal(n) * 2})
    out["sum_deadzone"] = (simple("SUM_DEADZONE", n, alpha=1.5, zone_m=0.5), 0.9, {"x": rng.standard_normal(n) * 2})
    out["sum_quantile_vec"] = (simple("SUM_QUANTILE", n, quantile_alpha=rng.random(n),
                                      quantile_beta=rng.random(n)), 1.1, {"x": rng.standard_normal(n) * 2})
    out["sum_quantile_scalar"] = (simple("SUM_QUANTILE", n, quantile_alpha=np.array([0.3]),
                                         quantile_beta=np.array([0.7])), 2.0, {"x": rng.standard_normal(n) * 2})
    out["non_negative"] = (simple("NON_NEGATIVE", n), 1.0, {"x": rng.standard_normal(n)})
    out["norm_2_shrink"] = (simple("NORM_2", 300, alpha=2.0), 1.5, {"x": rng.standard_normal(300)})
    out["norm_2_zero"] = (simple("NORM_2", 50, alpha=100.0), 1.0, {"x": rng.standard_normal(50)})
    out["sum_logistic"] = (simple("SUM_LOGISTIC", n), 1.7, {"x": rng.standard_normal(n) * 4})
    out["sum_logistic_large_lambda"] = (simple("SUM_LOGISTIC", 500, alpha=3.0), 25.0,
                                        {"x": rng.standard_normal(500) * 10})
    out["sum_exp"] = (simple("SUM_EXP", n, alpha=0.8), 1.2, {"x": rng.standard_normal(n) * 3})
    out["sum_neg_entr"] = (simple("SUM_NEG_ENTR", n, alpha=1.3), 0.9, {"x": rng.standard_normal(n) * 3 + 2})
    out["sum_inv_pos"] = (simple("SUM_INV_POS", n, alpha=0.6), 1.4, {"x": rng.standard_normal(n) * 3 + 1})
    out["sum_neg_log"] = (simple("SUM_NEG_LOG", n, alpha=1.1), 0.7, {"x": rng.standard_normal(n) * 3})
    out["tv_1d"] = (simple("TOTAL_VARIATION_1D", n, alpha=2.0), 1.5,
                    {"x": np.repeat(rng.standard_normal(20) * 3, 100) + rng.standard_normal(n)})
    out["tv_1d_len2"] = (simple("TOTAL_VARIATION_1D", 2, alpha=0.1), 1.0, {"x": np.array([1.0, -2.0])})
    out["tv_1d_len1"] = (simple("TOTAL_VARIATION_1D", 1, alpha=0.1), 1.0, {"x": np.array([3.0])})
    # SUM_SQUARE with a dense argument (least squares), wide and tall
    for nm, (m, nn) in {"sum_square_wide": (30, 80), "sum_square_tall": (120, 40)}.items():
        a = rng.standard_normal((m, nn))
        b = rng.standard_normal(m)
        x = Variable("x", nn)
        f = ProxFunction("SUM_SQUARE", alpha=0.5, args=[Affine(m, [
            Term(x, LinearMapSpec.Dense(a, name="A")), Term(b, LinearMapSpec.Scalar(-1.0, m), name="b")])])
        out[nm] = (f, 0.7, {"x": rng.standard_normal(nn)})

    # ---- multi-argument / sort-based operators (own generator: the cases above stay as they are) ----
    rng2 = np.random.default_rng(4321)

    def epi(ptype, n, **kw):
        x, t = Variable("x", n), Variable("t", 1)
        return ProxFunction(ptype, args=[Affine(n, [Term(x)]), Affine(1, [Term(t)])], epigraph=True, **kw)

    n = 1500
    out["max"] = (simple("MAX", n, alpha=0.9), 1.7, {"x": rng2.standard_normal(n) * 2})
    out["max_scaled_arg"] = (simple("MAX", n, alpha=0.5, scale=-1.5), 0.8, {"x": rng2.standard_normal(n)})
    out["max_epigraph"] = (epi("MAX", n), 1.0, {"x": rng2.standard_normal(n) * 2, "t": np.array([0.3])})
    out["max_epigraph_inside"] = (epi("MAX", 200), 1.0, {"x": rng2.standard_normal(200), "t": np.array([10.0])})
    out["sum_largest"] = (simple("SUM_LARGEST", n, alpha=0.8, sum_largest_k=7), 1.3,
                          {"x": rng2.standard_normal(n) * 2})
    out["sum_square_epigraph"] = (epi("SUM_SQUARE", n), 1.0, {"x": rng2.standard_normal(n), "t": np.array([3.0])})
    out["sum_square_epigraph_inside"] = (epi("SUM_SQUARE", 20), 1.0,
                                         {"x": rng2.standard_normal(20) * 0.1, "t": np.array([30.0])})
    out["norm_1_epigraph"] = (epi("NORM_1", n), 1.0, {"x": rng2.standard_normal(n) * 2, "t": np.array([5.0])})
    out["norm_1_epigraph_inside"] = (epi("NORM_1", 100), 1.0, {"x": rng2.standard_normal(100) * 0.01,
                                                             "t": np.array([5.0])})
    out["sum_hinge_epigraph"] = (epi("SUM_HINGE", n), 1.0, {"x": rng2.standard_normal(n) * 2, "t": np.array([2.0])})
    out["sum_deadzone_epigraph"] = (epi("SUM_DEADZONE", n, zone_m=0.5), 1.0,
                                    {"x": rng2.standard_normal(n) * 2, "t": np.array([1.0])})
    out["sum_quantile_epigraph"] = (epi("SUM_QUANTILE", n, quantile_alpha=np.array([0.3]),
                                        quantile_beta=np.array([0.7])), 1.0,
                                    {"x": rng2.standard_normal(n) * 2, "t": np.array([4.0])})
    xk, yk = Variable("x", 800), Variable("y", 800)
    out["sum_kl_div"] = (ProxFunction("SUM_KL_DIV", alpha=1.2, args=[Affine(800, [Term(xk)]), Affine(800, [Term(yk)])]),
                         0.8, {"x": rng2.random(800) * 3 + 0.1, "y": rng2.random(800) * 3 + 0.1})
    # NewtonEpigraph (newton.cc:108-185)
    n = 1000
    out["sum_exp_epigraph"] = (epi("SUM_EXP", n), 1.0, {"x": rng2.standard_normal(n), "t": np.array([300.0])})
    out["sum_exp_epigraph_inside"] = (epi("SUM_EXP", 50), 1.0, {"x": rng2.standard_normal(50) - 3,
                                                               "t": np.array([100.0])})
    out["sum_logistic_epigraph"] = (epi("SUM_LOGISTIC", n), 1.0, {"x": rng2.standard_normal(n) * 2,
                                                                 "t": np.array([400.0])})
    out["sum_inv_pos_epigraph"] = (epi("SUM_INV_POS", n), 1.0, {"x": rng2.random(n) * 2 + 0.2,
                                                               "t": np.array([500.0])})
    out["sum_inv_pos_epigraph_infeasible"] = (epi("SUM_INV_POS", 100), 1.0, {"x": rng2.standard_normal(100),
                                                                            "t": np.array([50.0])})

    # SECOND_ORDER_CONE (second_order_cone.cc): rows of [t X]
    def soc(m, n, at=1.0, ax=1.0, gt=None, gx=None):
        t, x = Variable("t", m), Variable("x", m * n)
        a0 = [Term(t) if at == 1.0 else Term(t, LinearMapSpec.Scalar(at, m))]
        a1 = [Term(x) if ax == 1.0 else Term(x, LinearMapSpec.Scalar(ax, m * n))]
        if gt is not None:
            a0.append(Term(gt, name="gt"))
        if gx is not None:
            a1.append(Term(gx, name="gx"))
        return ProxFunction("SECOND_ORDER_CONE", args=[Affine(m, a0), Affine(m * n, a1)], soc_dims=(m, n))

    out["soc_rows"] = (soc(300, 7), 1.0, {"t": rng2.standard_normal(300), "x": rng2.standard_normal(2100)})
    out["soc_scaled_shifted"] = (soc(40, 5, at=2.0, ax=-0.5, gt=rng2.standard_normal(40),
                                     gx=rng2.standard_normal(200)), 0.7,
                                 {"t": rng2.standard_normal(40), "x": rng2.standard_normal(200)})
    out["soc_single_cone"] = (soc(1, 500), 1.0, {"t": np.array([0.5]), "x": rng2.standard_normal(500)})
    xe, te = Variable("x", 700), Variable("t", 700)
    out["exp_epigraph"] = (ProxFunction("EXP", epigraph=True, args=[Affine(700, [Term(xe)]), Affine(700, [Term(te)])]),
                           1.0, {"x": rng2.standard_normal(700) * 2, "t": rng2.standard_normal(700) * 2 + 1})
    # implicit epigraph operators: Newton / bisection on the multiplier around the plain prox
    rng3 = np.random.default_rng(9876)
    n = 900
    out["sum_neg_log_epigraph"] = (epi("SUM_NEG_LOG", n), 1.0, {"x": rng3.standard_normal(n) * 2,
                                                                "t": np.array([-200.0])})
    out["sum_neg_log_epigraph_loose"] = (epi("SUM_NEG_LOG", 60), 1.0, {"x": rng3.random(60) * 3 + 0.5,
                                                                       "t": np.array([40.0])})
    out["sum_neg_entr_epigraph"] = (epi("SUM_NEG_ENTR", n), 1.0, {"x": rng3.random(n) * 3 + 0.05,
                                                                  "t": np.array([100.0])})
    out["sum_neg_entr_epigraph_inside"] = (epi("SUM_NEG_ENTR", 80), 1.0, {"x": rng3.random(80) + 0.1,
                                                                          "t": np.array([50.0])})
    xk, yk, tk = Variable("x", 600), Variable("y", 600), Variable("t", 1)
    out["sum_kl_div_epigraph"] = (ProxFunction("SUM_KL_DIV", epigraph=True, args=[
        Affine(600, [Term(xk)]), Affine(600, [Term(yk)]), Affine(1, [Term(tk)])]), 1.0,
        {"x": rng3.random(600) * 3 + 0.1, "y": rng3.random(600) * 3 + 0.1, "t": np.array([20.0])})
    out["log_sum_exp"] = (simple("LOG_SUM_EXP", n, alpha=1.3), 0.9, {"x": rng3.standard_normal(n) * 3})
    out["log_sum_exp_scaled_arg"] = (simple("LOG_SUM_EXP", 300, alpha=0.4, scale=2.0), 1.1,
                                     {"x": rng3.standard_normal(300)})
    out["log_sum_exp_epigraph"] = (epi("LOG_SUM_EXP", n), 1.0, {"x": rng3.standard_normal(n) * 2,
                                                                "t": np.array([3.0])})
    out["log_sum_exp_epigraph_inside"] = (epi("LOG_SUM_EXP", 40), 1.0, {"x": rng3.standard_normal(40),
                                                                        "t": np.array([30.0])})
    out["sum_largest_epigraph"] = (epi("SUM_LARGEST", n, sum_largest_k=7), 1.0,
                                   {"x": rng3.standard_normal(n) * 2, "t": np.array([5.0])})
    # orthogonally invariant matrix functions (ortho_invariant.cc): eigen-decomposition + a vector operator
    rng4 = np.random.default_rng(2468)

    def mat(ptype, m, n, epigraph=False, alpha=1.0):
        X = Variable("x", m * n)
        args = [Affine(m * n, [Term(X)])]
        if epigraph:
            args.append(Affine(1, [Term(Variable("t", 1))]))
        return ProxFunction(ptype, args=args, alpha=alpha, epigraph=epigraph, mat_dims=(m, n))

    def sym(n, shift=0.0):
        a = rng4.standard_normal((n, n))
        return ((a + a.T) / 2 + shift * np.eye(n)).ravel(order="F")

    out["lambda_max"] = (mat("LAMBDA_MAX", 12, 12, alpha=1.5), 0.8, {"x": sym(12)})
    out["lambda_max_epigraph"] = (mat("LAMBDA_MAX", 10, 10, epigraph=True), 1.0,
                                  {"x": sym(10), "t": np.array([0.5])})
    out["neg_log_det"] = (mat("NEG_LOG_DET", 15, 15, alpha=0.7), 1.2, {"x": sym(15, 1.0)})
    out["neg_log_det_epigraph"] = (mat("NEG_LOG_DET", 9, 9, epigraph=True), 1.0,
                                   {"x": sym(9, 2.0), "t": np.array([-20.0])})
    out["norm_nuclear_tall"] = (mat("NORM_NUCLEAR", 20, 8, alpha=1.1), 0.9,
                                {"x": rng4.standard_normal(160)})
    # (a 6 x 14 case was drawn here: for m < n the reference sizes its nested prox min(m, n) while the spectrum
    # of Y'Y has n entries, ortho_invariant.cc:77 vs :41-46, and returns its input; not a fixture)
    rng4.standard_normal(84)
    out["norm_nuclear_epigraph"] = (mat("NORM_NUCLEAR", 16, 11, epigraph=True), 1.0,
                                    {"x": rng4.standard_normal(176), "t": np.array([8.0])})
    out["semidefinite"] = (mat("SEMIDEFINITE", 16, 16), 1.0, {"x": rng4.standard_normal(256)})
    out["semidefinite_64"] = (mat("SEMIDEFINITE", 64, 64), 1.0, {"x": rng4.standard_normal(4096)})
    # ---- VectorProx::InitDiagonal (diagonal H, zeros included) and has_axis (own generator again) ----
    rng5 = np.random.default_rng(1357)

    def diag_arg(ptype, n, alpha=1.0, **kw):
        w = rng5.standard_normal(n)
        w[rng5.random(n) < 0.15] = 0.0
        x = Variable("x", n)
        return ProxFunction(ptype, args=[Affine(n, [Term(x, LinearMapSpec.Diagonal(w, name="w"))])], alpha=alpha, **kw)

    n = 1201
    out["norm_1_diag"] = (diag_arg("NORM_1", n, alpha=0.8), 1.4, {"x": rng5.standard_normal(n) * 2})
    out["sum_hinge_diag"] = (diag_arg("SUM_HINGE", n), 0.9, {"x": rng5.standard_normal(n) * 2})
    out["sum_deadzone_diag"] = (diag_arg("SUM_DEADZONE", n, alpha=1.2, zone_m=0.4), 0.7, {"x": rng5.standard_normal(n) * 2})
    out["non_negative_diag"] = (diag_arg("NON_NEGATIVE", n), 1.0, {"x": rng5.standard_normal(n)})
    out["sum_logistic_diag"] = (diag_arg("SUM_LOGISTIC", 400), 1.3, {"x": rng5.standard_normal(400) * 3})

    def axis_norm2(m, k, axis, alpha):
        x = Variable("x", m * k)
        return ProxFunction("NORM_2", args=[Affine(m * k, [Term(x)])], alpha=alpha, axis=axis, axis_dims=(m, k))

    out["norm_2_axis0"] = (axis_norm2(37, 11, 0, 2.5), 1.2, {"x": rng5.standard_normal(37 * 11)})
    out["norm_2_axis1"] = (axis_norm2(37, 11, 1, 1.5), 0.9, {"x": rng5.standard_normal(37 * 11)})
    out["norm_2_axis1_wide"] = (axis_norm2(5, 1500, 1, 20.0), 1.0, {"x": rng5.standard_normal(5 * 1500)})
    return out


def solve_case(name, ref_runner):
    prob, info, params = make_problem(name)
    r = ref_runner.run(prob, params=params, mode="trace")
    stock = ref_runner.run(prob, params=params, mode="solve")
    # the stepped run must end exactly where the stock Solve() ends
    for k in TRACE_KEYS:
        assert r["status"][k] == stock["status"][k], (name, k, r["status"], stock["status"])
    trace = np.array([[t[k] for k in TRACE_KEYS] for t in r["trace"]], dtype=np.float64)
    v = info["solution_var"]
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"), trace=trace,
        status=np.array([r["status"][k] for k in TRACE_KEYS], dtype=np.float64),
        mnN=np.array([r["m"], r["n"], r["N"]], dtype=np.int64),
        x=r["x"][v].astype(np.float64), objective=np.float64(info["objective"](r["x"])))
    return {"digest": problem_digest(prob), "solution_var": v,
                            "epochs": int(trace.shape[0]),
                            "num_iterations": int(r["status"]["num_iterations"]),
                            "state": int(r["status"]["state"])}


def add_missing_prox():
    """Run the reference only on the single-operator cases that have no fixture yet."""
    from oracle import ref_runner
    assert ref_runner.available(), "build oracle/_ref first: make -C oracle"
    with open(os.path.join(HERE, "index.json")) as fh:
        index = json.load(fh)
    for name in SOLVE_CASES:
        if name in index["solve"]:
            continue
        index["solve"][name] = solve_case(name, ref_runner)
        print(name, index["solve"][name])
    for name, (f, lam, v) in prox_cases().items():
        if name in index["prox"]:
            continue
        x = ref_runner.run_prox(f, lam, v)
        np.savez_compressed(os.path.join(HERE, "prox_" + name + ".npz"), lam=np.float64(lam),
                            **{"v_" + k: val for k, val in v.items()},
                            **{"x_" + k: val for k, val in x.items() if k in v})
        index["prox"][name] = {"type": f.type, "n": int(sum(val.size for val in v.values()))}
        print("prox", name, index["prox"][name])
    with open(os.path.join(HERE, "index.json"), "w") as fh:
        json.dump(index, fh, indent=1, sort_keys=True)


def main():
    from oracle import ref_runner
    assert ref_runner.available(), "build oracle/_ref first: make -C oracle"
    index = {"solve": {}, "prox": {}}
    for name in SOLVE_CASES:
        index["solve"][name] = solve_case(name, ref_runner)
        print(name, index["solve"][name])
    # two-block solver: the reference offers no stepping hook for it, so the per-epoch
    # residuals come from stock Solve() runs cut at max_iterations = k+1 (the final
    # ComputeResiduals of such a run sees exactly the state of epoch k, prox_admm_two_block.cc:123-126)
    index["two_block"] = {}
    for name in TWO_BLOCK_CASES:
        prob, info, params = make_problem(name)
        p = dict(params, solver="PROX_ADMM_TWO_BLOCK")
        final = ref_runner.run(prob, params=p, mode="solve")
        fs = final["status"]
        rows = []
        last_epoch = fs["num_iterations"] if fs["state"] == 3 else (fs["num_iterations"] - 1) // 10 * 10
        for k in range(0, last_epoch + 1, 10):
            if fs["state"] == 3 and k == fs["num_iterations"]:
                break
            cut = ref_runner.run(prob, params=dict(p, max_iterations=k + 1), mode="solve")["status"]
            assert cut["num_iterations"] == k + 1 and cut["state"] == 4
            rows.append([2, k] + [cut[x] for x in TRACE_KEYS[2:]])
        rows.append([fs[x] for x in TRACE_KEYS])
        v = info["solution_var"]
        np.savez_compressed(os.path.join(HERE, name + ".npz"), trace=np.array(rows, dtype=np.float64),
                            status=np.array([fs[x] for x in TRACE_KEYS], dtype=np.float64),
                            x=final["x"][v].astype(np.float64),
                            objective=np.float64(info["objective"](final["x"])))
        index["two_block"][name] = {"digest": problem_digest(prob), "solution_var": v, "epochs": len(rows),
                                    "num_iterations": int(fs["num_iterations"]), "state": int(fs["state"])}
        print(name, index["two_block"][name])
    for name, (f, lam, v) in prox_cases().items():
        x = ref_runner.run_prox(f, lam, v)
        np.savez_compressed(os.path.join(HERE, "prox_" + name + ".npz"), lam=np.float64(lam),
                            **{"v_" + k: val for k, val in v.items()},
                            **{"x_" + k: val for k, val in x.items() if k in v})
        index["prox"][name] = {"type": f.type, "n": int(sum(val.size for val in v.values()))}
        print("prox", name, index["prox"][name])
    with open(os.path.join(HERE, "index.json"), "w") as fh:
        json.dump(index, fh, indent=1, sort_keys=True)


if __name__ == "__main__":
    if "--missing-prox" in sys.argv:
        add_missing_prox()
    else:
        main()
