"""CPU restatement (numpy / scipy) of the reference's ProxADMM hot path.

TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module; the product (epsilon_b200/) never does, and
it fails loudly when its CUDA library is missing.

What is restated (reference = mfouda/epsilon, paths relative to /root/reference):
  * ProxADMMSolver::{Init,Solve,ComputeResiduals,GetSolution}  src/epsilon/algorithms/prox_admm.cc:25-217
  * VectorProx scalar / diagonal reduction                      src/epsilon/prox/vector_prox.cc:51-183
  * ScaledZoneProx (NORM_1, SUM_HINGE, SUM_DEADZONE, SUM_QUANTILE) src/epsilon/prox/scaled_zone.cc:34-104
  * NonNegativeProx, Norm2Prox                                  src/epsilon/prox/non_negative.cc:5-9, norm_2.cc:6-17
  * NewtonProx + Logistic (global line search)                  src/epsilon/prox/newton.cc:8-14,47-91, sum_logistic.cc:8-32
  * TotalVariation1DProx -> Johnson's dynamic program           src/epsilon/prox/total_variation_1d.cc:12-24,
                                                                third_party/glmgen/c_lib/glmgen/src/tf/tf_dp.c:43-170
  * SumSquareProx / AffineProx (KKT solve by block elimination) src/epsilon/prox/sum_square.cc:10-40, affine.cc:19-49,
                                                                src/epsilon/vector/block_cholesky.cc:119-137
  * EvalProx                                                    python/epopt/solvemodule.cc:189-242

Parity status: PINNED.  The reference's own tests hold no golden vectors for this path
(SURVEY.md section 8c), so this restatement is pinned against outputs of the reference
itself: the unmodified reference solver compiled by oracle/Makefile (oracle/_ref/epsilon_ref)
was run in the build container on seeded problems by tests/golden/make_golden.py and its
per-epoch residual traces, iteration counts, solutions and single-operator outputs are
committed under tests/golden/; tests/test_oracle_golden.py holds this module to them.

Data layout: the reference's BlockVector (std::map<string, VectorXd>, missing key = 0,
src/epsilon/vector/block_vector.cc:43-48) is a flat vector here, constraint rows
concatenated in constraint order; "missing = 0" is then automatic.  Sums over keys
(BlockVector::norm, block_vector.cc:87-93) therefore run in a different order than the
reference's; the difference is rounding only (~1e-16 relative).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Dict, List, Optional

import numpy as np
import scipy.linalg as sla
import scipy.sparse as sp

HERE = os.path.dirname(os.path.abspath(__file__))

STATE_RUNNING, STATE_OPTIMAL, STATE_MAX_ITERATIONS = 2, 3, 4

# SolverParams defaults, proto/epsilon/solver_params.proto:4-71
DEFAULT_PARAMS = {"max_iterations": 10000, "rho": 1.0, "abs_tol": 1e-4, "rel_tol": 1e-2,
                  "epoch_iterations": 10}

KKT_TYPES = ("SUM_SQUARE", "AFFINE", "CONSTANT")
ZONE_TYPES = ("NORM_1", "SUM_HINGE", "SUM_DEADZONE", "SUM_QUANTILE")


# ---------------------------------------------------------------------------
# elementwise / block proximal operators (ApplyVector of each VectorProx subclass)
# ---------------------------------------------------------------------------

def scaled_zone(v: np.ndarray, lam: np.ndarray, alpha, beta, M: float, C: float = 0.0) -> np.ndarray:
    """ApplyScaledZoneProx, scaled_zone.cc:78-104.  Branch order as written there; the
    products lam*alpha are rounded before the add (the reference build has no FMA)."""
    x = v - C
    la = lam * alpha
    lb = lam * beta
    out = np.where(x > 0, M, -M) * np.ones_like(x)        # :97-100
    lo = x < (-M - lb)                                      # :95-96
    out = np.where(lo, x + lb, out)
    hi = x > (M + la)                                       # :93-94 (tested before :95)
    out = np.where(hi, x - la, out)
    inside = np.abs(x) <= M                                 # :91-92 (tested first)
    out = np.where(inside, x, out)
    return out


def zone_params(f, n: int):
    """GetParams, scaled_zone.cc:34-76 -> (alpha, beta, M)."""
    if f.type == "NORM_1":
        return 1.0, 1.0, 0.0
    if f.type == "SUM_DEADZONE":
        return 1.0, 1.0, float(f.zone_m)
    if f.type == "SUM_HINGE":
        return 1.0, 0.0, 0.0
    if f.type == "SUM_QUANTILE":
        def promote(a):  # scaled_zone.cc:26-32
            a = np.asarray(a, dtype=np.float64).ravel()
            if a.size == n:
                return a
            assert a.size == 1, "%d != %d" % (a.size, n)
            return np.full(n, a[0])
        return promote(f.quantile_alpha), promote(f.quantile_beta), 0.0
    raise ValueError(f.type)


def norm2_block(v: np.ndarray, lam: float) -> np.ndarray:
    """Norm2Prox::ApplyVector, norm_2.cc:6-17."""
    nv = float(np.sqrt(v @ v))
    if nv >= lam:
        return (1 - lam / nv) * v
    return np.zeros_like(v)


def _logistic_grad(x):   # sum_logistic.cc:18-24 (unguarded exp, as written)
    with np.errstate(over="ignore", invalid="ignore"):
        e = np.exp(x)
        return e / (1 + e)


def _logistic_hess(x):   # sum_logistic.cc:25-31
    with np.errstate(over="ignore", invalid="ignore"):
        e = np.exp(x)
        return e / (1 + e) ** 2


def newton_logistic(v: np.ndarray, lam: np.ndarray) -> np.ndarray:
    """ApplyNewtonProx with f = Logistic, newton.cc:47-91: damped Newton on
    x - v + lam*sigma(x) = 0 with backtracking on the WHOLE-vector residual norm."""
    n = v.size
    eps = max(1e-12, 1e-10 / n)                                 # :52
    x = v.copy()                                                # :55 proj_feasible = identity
    for _ in range(100):                                        # :58-59
        gx = x - v + lam * _logistic_grad(x)                    # ProxResidual, newton.cc:8-14
        dx = gx / (1.0 + lam * _logistic_hess(x))               # hess_inv, newton.h:26-31
        theta, x_res = 1.0, float(np.sqrt(gx @ gx))             # :65-68
        while theta > eps:                                      # :69-79
            nx = x - theta * dx
            r = nx - v + lam * _logistic_grad(nx)
            nx_res = float(np.sqrt(r @ r))
            if nx_res <= (1 - 0.001 * theta) * x_res:
                x, x_res = nx, nx_res
                break
            theta *= 0.5
        if x_res < eps:                                         # :81-83
            break
    return x


def newton_prox(v: np.ndarray, lam: np.ndarray, gradf, hessf, proj=lambda x: x) -> np.ndarray:
    """ApplyNewtonProx (newton.cc:47-91) for any ElemwiseSmoothFunction (newton.h:23-32)."""
    n = v.size
    eps = max(1e-12, 1e-10 / n)
    x = proj(v.copy())
    with np.errstate(all="ignore"):
        for _ in range(100):
            gx = x - v + lam * gradf(x)
            dx = gx / (1.0 + lam * hessf(x))
            theta, x_res = 1.0, float(np.sqrt(gx @ gx))
            while theta > eps:
                nx = proj(x - theta * dx)
                r = nx - v + lam * gradf(nx)
                nx_res = float(np.sqrt(r @ r))
                if nx_res <= (1 - 0.001 * theta) * x_res:
                    x, x_res = nx, nx_res
                    break
                theta *= 0.5
            if x_res < eps:
                break
    return x


_FEASIBLE = lambda x: np.maximum(x, 1e-6)   # proj_feasible, sum_neg_entr.cc:40-42, sum_inv_pos.cc:37-39
SMOOTH = {
    # type: (gradf, hessf, proj_feasible)
    "SUM_EXP": (np.exp, np.exp, lambda x: x),                                        # sum_exp.cc:10-36
    "SUM_NEG_ENTR": (lambda x: 1 + np.log(x), lambda x: 1 / x, _FEASIBLE),           # sum_neg_entr.cc:10-42
    "SUM_INV_POS": (lambda x: -1 / (x * x), lambda x: 2 / (x * x * x), _FEASIBLE),   # sum_inv_pos.cc:10-41
}


def neg_log_prox(v: np.ndarray, lam: np.ndarray) -> np.ndarray:
    """ApplyNegLogProx, sum_neg_log.cc:10-26 (closed form)."""
    s = np.sqrt(v * v + 4 * lam)
    with np.errstate(all="ignore"):
        return np.where(v >= 0, (v + s) / 2, 2 * lam / (-v + s))



# ---------------------------------------------------------------------------
# sort-based and epigraph operators (stacked arguments: v = [arg0 | arg1 | ...])
# ---------------------------------------------------------------------------

def max_prox(v: np.ndarray, lam: float) -> np.ndarray:
    """MaxProx::ApplyVector, max.cc:8-41: x_i = min(v_i, t), sum_i (v_i - t)_+ = lam."""
    y = np.sort(v)[::-1]
    acc = np.cumsum(np.concatenate([[-lam], y]))           # acc before step i, in the reference's order
    i = np.arange(y.size)
    stop = np.nonzero(y * i < acc[:-1])[0]
    k = int(stop[0]) if stop.size else y.size
    t = acc[k] / k if k > 0 else 0.0
    return np.minimum(v, t)


def max_epigraph(v: np.ndarray, s: float):
    """MaxEpigraph::ApplyVector, max.cc:46-85: projection onto {max_i x_i <= t}."""
    y = np.sort(v)[::-1]
    if s >= y[0]:
        return v.copy(), s
    acc = np.cumsum(np.concatenate([[0.0], y - s]))
    div = 1.0 + np.arange(y.size)
    stop = np.nonzero(div * (y - s) < acc[:-1])[0]
    k = int(stop[0]) if stop.size else y.size
    delta = acc[k] / (1.0 + k) if k > 0 else 0.0
    t = s + delta
    return v - np.maximum(0.0, v - t), t


def sum_largest_prox(v: np.ndarray, lam: float, k: int) -> np.ndarray:
    """SumLargestProx::ApplyVector, sum_largest.cc:17-58 (two-pointer sweep, as written)."""
    y = np.sort(v)[::-1]
    n = y.size
    q, acc, inside, i, j = 0.0, -k * lam, 0, 0, 0
    while i < n and j < n:
        if y[i] * inside <= acc and (y[j] - lam) * inside <= acc:
            break
        if y[i] >= y[j] - lam:
            acc += y[i]; inside += 1; i += 1
        else:
            acc += -y[j] + lam; inside -= 1; j += 1
        q = acc / inside if inside else np.copysign(np.inf, acc) if acc else np.nan
    return v - np.maximum(0.0, np.minimum(lam, v - q))


def largest_real_cubic_root(b: float, c: float, d: float) -> float:
    """LargestRealCubicRoot, newton.cc:293-323 (Durand-Kerner from the reference's start)."""
    eps = 1e-12
    p = complex(0.4, 0.9); q = p * p; r = p * p * p
    for _ in range(100):
        fp = p * p * p + b * p * p + c * p + d
        np_ = p - fp / ((p - q) * (p - r))
        fq = q * q * q + b * q * q + c * q + d
        nq = q - fq / ((q - p) * (q - r))
        fr = r * r * r + b * r * r + c * r + d
        nr = r - fr / ((r - p) * (r - q))
        if abs(fp) < eps and abs(fq) < eps and abs(fr) < eps:
            break
        p, q, r = np_, nq, nr
    m = -1e41
    for z in (p, q, r):
        if abs(z.imag) < eps and z.real > m:
            m = z.real
    return m


def sum_square_epigraph(u: np.ndarray, s: float):
    """SumSquareEpigraph::ApplyVector, sum_square.cc:42-57."""
    lam = largest_real_cubic_root(1 + s, 0.25 + s, (s - float(u @ u)) / 4)
    lam = max(lam, 0.0)
    return u / (1 + 2 * lam), s + lam


def scaled_zone_epigraph(v: np.ndarray, s: float, alpha: float, beta: float, M: float, C: float = 0.0):
    """ScaledZoneEpigraph::ApplyVector, scaled_zone.cc:152-279, for constant alpha/beta (the
    reference indexes alpha/beta by the position in its permuted work array, :238-252, which is
    only meaningful when they are constant).  The reference finds lam by a randomised
    three-way-partition select; the root it brackets is the unique solution of
        sum_{|key_k| > lam} w_k^2 (|key_k| - lam) = s + lam,   lam = acc / (div + 1)  (:268)
    which is computed here from the sorted keys; the two differ by summation order only."""
    y = v - C
    keep = ~((np.abs(y) <= M) | ((y > 0) & (alpha == 0)) | ((y < 0) & (beta == 0)))
    yk = y[keep]
    fval = float(np.sum(np.where(yk > M, alpha * (yk - M), beta * (-yk - M))))
    if fval <= s:
        return v.copy(), s
    w2 = np.where(yk > 0, alpha * alpha, beta * beta)
    key = np.where(yk > 0, (yk - M) / alpha if alpha else 0.0, -(yk + M) / beta if beta else 0.0)  # |key|
    order = np.argsort(-key, kind="stable")
    key, w2 = key[order], w2[order]
    acc = -s + np.cumsum(w2 * key)
    div = np.cumsum(w2)
    lam_k = acc / (div + 1)
    nxt = np.concatenate([key[1:], [0.0]])
    ok = np.nonzero(lam_k >= nxt)[0]
    lam = float(lam_k[ok[0]])
    return scaled_zone(v, np.full(v.size, lam), alpha, beta, M, C), s + lam


def kl_div_prox(lam: np.ndarray, u: np.ndarray, v: np.ndarray):
    """ApplyKLDivProx, sum_kl_div.cc:6-52 (per-element Newton on yhat, as written)."""
    eps = 1e-13
    n = u.size
    x, y = np.empty(n), np.empty(n)
    for i in range(n):
        li, ui, vi = float(lam[i]), float(u[i]), float(v[i])
        yhat = max((0.5 + li - vi) / li, eps)
        if abs(ui) < eps * eps and abs(vi) < eps * eps:
            x[i], y[i] = ui, vi
            continue
        for _ in range(1000):
            f = li * yhat * yhat + (vi - li) * yhat - ui + li * np.log(yhat)
            F = 2 * li * yhat + (vi - li) + li / yhat
            res = f / F
            if abs(res) < eps or (yhat <= eps * 2 and res > 0) or (li * yhat + vi - li <= eps * 2 and res > 0):
                break
            yhat = yhat - res
            if yhat < eps:
                yhat = eps
            if li * yhat + vi - li < eps:
                yhat = (eps + li - vi) / li
        y[i] = li * yhat + vi - li
        x[i] = y[i] * yhat
    return x, y


# eval of the ElemwiseSmoothFunction classes (sum over the elements)
SMOOTH_EVAL = {
    "SUM_EXP": lambda x: float(np.sum(np.exp(x))),                       # sum_exp.cc:12-19
    "SUM_LOGISTIC": lambda x: float(np.sum(np.log(1 + np.exp(x)))),      # sum_logistic.cc:10-17
    "SUM_INV_POS": lambda x: float(np.sum(1 / x)),                       # sum_inv_pos.cc:12-19
}


def newton_epigraph(v: np.ndarray, s: float, evalf, gradf, hessf, proj=lambda x: x, hinv=None):
    """NewtonEpigraph::ApplyVector, newton.cc:108-185: Newton on (x, t, lam) with the arrowhead
    structure of the KKT system and a backtracking line search on the residual norm."""
    n = v.size
    eps = max(1e-12, 1e-10 / n)

    def residual(lam, x, t):                                            # EpigraphResidual, newton.cc:16-27
        r = np.empty(n + 2)
        r[:n] = x - v + lam * gradf(x)
        r[n] = t - s - lam
        r[n + 1] = evalf(x) - t
        return r

    with np.errstate(all="ignore"):
        x = proj(v.copy())
        if float(np.linalg.norm(v - x)) < eps and evalf(x) <= s:
            return v.copy(), s
        t, lam = s, 1.0
        for _ in range(100):
            g = residual(lam, x, t)
            nt_step = hinv(lam, x, g[:n]) if hinv else g[:n] / (1.0 + lam * hessf(x))
            nt_res = float(g[:n] @ nt_step)
            scale = (-nt_res + g[n] + g[n + 1]) / (nt_res + 1)
            step_x = (1 + scale) * nt_step
            step_t = g[n] - scale
            step_l = -scale
            theta, x_res = 1.0, float(np.linalg.norm(g))
            while theta > eps:
                nx = proj(x - theta * step_x)
                nt = t - theta * step_t
                nlam = lam - theta * step_l
                if nlam < eps:
                    nlam = eps
                nx_res = float(np.linalg.norm(residual(nlam, nx, nt)))
                if nx_res <= (1 - 0.001 * theta) * x_res:
                    x, t, lam, x_res = nx, nt, nlam, nx_res
                    break
                theta *= 0.5
            if x_res < eps:
                break
    return x, t


def exp_epigraph(v: np.ndarray, s: np.ndarray):
    """ExpEpigraph::ApplyVector, exp.cc:12-78: elementwise projection onto {exp(x_i) <= t_i},
    Newton on (x, t, l) for all elements at once until every residual is below 1e-12."""
    x, t, l = v.copy(), s.copy(), np.ones(v.size)
    eps = 1e-12
    with np.errstate(all="ignore"):
        for _ in range(100):
            ex = np.exp(x)
            r0 = (x - v) + l * ex
            r1 = t - s - l
            r2 = ex - t
            if np.abs(r0).max() < eps and np.abs(r1).max() < eps and np.abs(r2).max() < eps:
                break
            h = 1 + l * ex
            d = ex
            dl = (-d * r0 + h * (r1 + r2)) / (d * d + h)
            dx = -(d * dl + r0) / h
            dt = dl - r1
            x, t, l = x + dx, t + dt, l + dl
        easy = np.exp(v) <= s
    x[easy], t[easy] = v[easy], s[easy]
    return x, t


def neg_log_epigraph(v: np.ndarray, s: float):
    """SumNegLogEpigraph::ApplyVector, sum_neg_log.cc:41-88: scalar Newton on the multiplier lambda of
    -sum log x_i(lambda) - lambda - s = 0 with x(lambda) the closed-form prox; no easy case."""
    lam = 1.0
    with np.errstate(all="ignore"):
        for _ in range(1000):
            z = np.sqrt(v * v + 4 * lam)
            pos = v >= 0
            gi = np.where(pos, -np.log((v + z) / 2), np.log((-v + z) / (2 * lam)))
            hi = np.where(pos, 1.0 / (v * (v + z) / 2 + 2 * lam), 1.0 / (v * 2 * lam / (-v + z) + 2 * lam))
            g, h = -lam - s, -1.0
            for i in range(v.size):                      # the reference accumulates element by element
                g += gi[i]; h -= hi[i]
            if abs(g) <= 1e-10:
                break
            if h >= -1e-10:
                h = -1e-10
            lam -= g / h
            if lam <= 1e-10:
                lam = 1e-10
    return neg_log_prox(v, np.full(v.size, lam)), s + lam


def implicit_newton_epigraph(v: np.ndarray, s: float, evalf, gradf, hessf, proj=lambda x: x):
    """ImplicitNewtonEpigraph::ApplyVector, newton.cc:192-238: Newton on lambda of f(prox_{lambda f}(v)) - lambda - s."""
    x = proj(v.copy())
    with np.errstate(all="ignore"):
        if evalf(x) <= s:
            return x, s
        lam = 1.0
        for _ in range(100):
            x = newton_prox(v, np.full(v.size, lam), gradf, hessf, proj)
            gx = gradf(x)
            glam = evalf(x) - lam - s
            hlam = -float((gx / (1.0 + lam * hessf(x))) @ gx) - 1
            if abs(glam) < 1e-10:
                break
            lam = lam - glam / hlam
            if lam < 0:
                lam = 1e-6
        return newton_prox(v, np.full(v.size, lam), gradf, hessf, proj), s + lam


def _neg_entr_eval(x):                                   # sum_neg_entr.cc:12-21 (skips x_i <= 0)
    with np.errstate(all="ignore"):
        return float(np.sum(np.where(x > 0, x * np.log(np.where(x > 0, x, 1.0)), 0.0)))


def kl_div_epigraph(u: np.ndarray, v: np.ndarray, s: float):
    """SumKLDivEpigraph::ApplyVector, sum_kl_div.cc:68-118."""
    eps, n, lam = 1e-10, u.size, 1.0
    with np.errstate(all="ignore"):
        for _ in range(100):
            x, y = kl_div_prox(np.full(n, lam), u, v)
            glam, hlam = -s - lam, -1.0
            for i in range(n):
                glam += x[i] * np.log(x[i] / y[i]) - x[i] + y[i]
                g0, g1 = np.log(x[i] / y[i]), -x[i] / y[i] + 1
                h00 = 1 + lam / x[i]; h01 = -lam / y[i]; h11 = 1 + lam * x[i] / (y[i] * y[i])
                det = h00 * h11 - h01 * h01
                hlam -= (g0 * (h11 * g0 - h01 * g1) + g1 * (h00 * g1 - h01 * g0)) / det
            if abs(glam) < eps or (lam <= eps * 2 and glam / hlam > 0):
                break
            lam = lam - glam / hlam
            if lam < eps:
                lam = eps
        x, y = kl_div_prox(np.full(n, lam), u, v)
    return x, y, s + lam


def _lse_eval(x):                                        # log_sum_exp.cc:22-30
    m = x.max()
    return float(m + np.log(np.exp(x - m).sum()))


def _lse_grad(x):                                        # log_sum_exp.cc:31-40
    w = np.exp(x - x.max())
    return w / w.sum()


def _lse_hess_inv(lam: float, x: np.ndarray, v: np.ndarray) -> np.ndarray:
    """LogSumExp::hess_inv, log_sum_exp.cc:41-64: (I + lam (diag(w) - w w'))^{-1} v; lam = lambda_(0)."""
    w = _lse_grad(x)
    t = r = 0.0
    for i in range(x.size):
        t += w[i] * w[i] / (1 + lam * w[i])
        r += v[i] * w[i] / (1 + lam * w[i])
    sc = lam * r / (1 - lam * t)
    return v / (1 + lam * w) + w / (1 + lam * w) * sc


def log_sum_exp_prox(v: np.ndarray, lam: float) -> np.ndarray:
    """ApplyNewtonProx (newton.cc:47-91) with the LogSumExp function (full Hessian through hess_inv)."""
    n = v.size
    eps = max(1e-12, 1e-10 / n)
    x = v.copy()
    for _ in range(100):
        gx = x - v + lam * _lse_grad(x)
        dx = _lse_hess_inv(lam, x, gx)
        theta, x_res = 1.0, float(np.linalg.norm(gx))
        while theta > eps:
            nx = x - theta * dx
            nx_res = float(np.linalg.norm(nx - v + lam * _lse_grad(nx)))
            if nx_res <= (1 - 0.001 * theta) * x_res:
                x, x_res = nx, nx_res
                break
            theta *= 0.5
        if x_res < eps:
            break
    return x


def log_sum_exp_epigraph(v: np.ndarray, s: float):
    """NewtonEpigraph::ApplyVector (newton.cc:108-185) with LogSumExp: the x block of the arrowhead
    step goes through LogSumExp::hess_inv."""
    return newton_epigraph(v, s, _lse_eval, _lse_grad, None, hinv=_lse_hess_inv)


def bisection_epigraph(v: np.ndarray, s: float, prox, evalf):
    """BisectionEpigraph::ApplyVector, newton.cc:240-286: doubling then bisection on lambda of
    f(prox_{lambda f}(v)) - lambda - s to |g| <= 1e-5.  (After 100 rounds without reaching it the
    reference leaves t unset; reported here as nan.)"""
    if evalf(v) <= s:
        return v.copy(), s
    lam, upper, lower, upper_fixed = 1.0, 1.0, 0.0, False
    x = v
    for _ in range(100):
        x = prox(v, lam)
        g = evalf(x) - (lam + s)
        if abs(g) <= 1e-5:
            return x, lam + s
        if g > 0 and not upper_fixed:
            lam *= 2
            upper = lam
        elif g > 0:
            lower = lam
            lam = (lam + upper) / 2
        else:
            upper = lam
            lam = (lam + lower) / 2
            upper_fixed = True
    return x, float("nan")


def _sum_largest_eval(x: np.ndarray, k: int) -> float:   # SumLargestProx::Eval, sum_largest.cc:60-73
    y = np.sort(x)[::-1]
    tot = 0.0
    for i in range(k):
        tot += y[i]
    return tot

# orthogonally invariant matrix functions: type -> (vector operator on the spectrum, symmetric_part, add_residual)
ORTHO_INVARIANT = {
    "LAMBDA_MAX": ("MAX", True, False),             # lambda_max.cc
    "NEG_LOG_DET": ("SUM_NEG_LOG", True, False),    # neg_log_det.cc
    "NORM_NUCLEAR": ("NORM_1", False, False),       # norm_nuclear.cc
    "SEMIDEFINITE": ("NON_NEGATIVE", True, True),   # semidefinite.cc
}


def ortho_invariant(ptype: str, epigraph: bool, m: int, n: int, lam: float, y: np.ndarray, s: float = 0.0):
    """OrthoInvariantProx::ApplyVector, ortho_invariant.cc:18-73.  The nested "eigen prox" is a VectorProx with
    H = I, A = (1/sqrt(lam)) I fed with d/sqrt(lam) (InitEigenProx :75-97, ApplyEigenProx :99-104), i.e.
    prox_{lam f}(d); in the epigraph form it is the plain projection of (d, s).  Eigen's
    SelfAdjointEigenSolver is LAPACK's eigh here: X = U f(d) V' does not depend on the eigenvector signs."""
    inner, symmetric_part, add_residual = ORTHO_INVARIANT[ptype]
    Y = y.reshape((m, n), order="F")
    if symmetric_part:
        d, V = np.linalg.eigh((Y + Y.T) / 2)
        U = V
    else:
        d, V = np.linalg.eigh(Y.T @ Y + 1e-15 * np.eye(n))
        d = np.sqrt(np.maximum(d, 0.0))
        dinv = np.where(d != 0, 1.0 / np.where(d != 0, d, 1.0), 0.0)
        U = Y @ V @ np.diag(dinv)
    t = None
    if epigraph:
        if inner == "MAX":
            xt, t = max_epigraph(d, s)
        elif inner == "SUM_NEG_LOG":
            xt, t = neg_log_epigraph(d, s)
        else:
            xt, t = scaled_zone_epigraph(d, s, 1.0, 1.0, 0.0)
    elif inner == "MAX":
        xt = max_prox(d, lam)
    elif inner == "SUM_NEG_LOG":
        xt = neg_log_prox(d, np.full(d.size, lam))
    elif inner == "NORM_1":
        xt = scaled_zone(d, np.full(d.size, lam), 1.0, 1.0, 0.0)
    else:
        xt = np.maximum(d, 0.0)
    X = U @ np.diag(xt) @ V.T
    if not epigraph:
        if add_residual:
            X = X + (Y - Y.T) / 2
        elif symmetric_part:
            X = (X + X.T) / 2
    return X.ravel(order="F"), t


# ---------------------------------------------------------------------------
# 1-D total variation (C restatement in oracle/tv1d_dp.c; pure Python for small n)
# ---------------------------------------------------------------------------

_LIB_PATH = os.path.join(HERE, "_build", "liboracle.so")
_lib = None


def build_c(force: bool = False) -> str:
    """gcc the C part of the oracle into oracle/_build/liboracle.so."""
    src = os.path.join(HERE, "tv1d_dp.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(_LIB_PATH), exist_ok=True)
        subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-o", _LIB_PATH, src],
                       check=True)
    return _LIB_PATH


def _c_lib():
    global _lib
    if _lib is None:
        try:
            lib = ctypes.CDLL(build_c())
            lib.oracle_tv1d.restype = ctypes.c_int
            lib.oracle_tv1d.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_double, ctypes.c_void_p]
            _lib = lib
        except Exception:
            _lib = False
    return _lib or None


def tv1d_py(y: np.ndarray, lam: float) -> np.ndarray:
    """Same dynamic program as oracle/tv1d_dp.c in plain Python (small n only)."""
    n = y.size
    if n <= 1 or lam == 0:
        return y.astype(np.float64).copy()
    pos, da, db = [0.0] * (2 * n), [0.0] * (2 * n), [0.0] * (2 * n)
    clo, chi = [0.0] * (n - 1), [0.0] * (n - 1)
    head, tail = n - 1, n
    clo[0], chi[0] = -lam + y[0], lam + y[0]
    pos[head], da[head], db[head] = clo[0], 1.0, -y[0] + lam
    pos[tail], da[tail], db[tail] = chi[0], -1.0, y[0] + lam
    al, bl, ar, br = 1.0, -y[1] - lam, -1.0, y[1] - lam
    for k in range(1, n - 1):
        a_lo, b_lo, lo = al, bl, head
        while lo <= tail:
            if a_lo * pos[lo] + b_lo > -lam:
                break
            a_lo += da[lo]; b_lo += db[lo]; lo += 1
        a_hi, b_hi, hi = ar, br, tail
        while hi >= lo:
            if -a_hi * pos[hi] - b_hi < lam:
                break
            a_hi += da[hi]; b_hi += db[hi]; hi -= 1
        clo[k] = (-lam - b_lo) / a_lo
        head = lo - 1
        pos[head] = clo[k]
        chi[k] = (lam + b_hi) / (-a_hi)
        tail = hi + 1
        pos[tail] = chi[k]
        da[head], db[head] = a_lo, b_lo + lam
        da[tail], db[tail] = a_hi, b_hi + lam
        al, bl, ar, br = 1.0, -y[k + 1] - lam, -1.0, y[k + 1] - lam
    a_lo, b_lo, lo = al, bl, head
    while lo <= tail:
        if a_lo * pos[lo] + b_lo > 0:
            break
        a_lo += da[lo]; b_lo += db[lo]; lo += 1
    out = np.empty(n)
    out[n - 1] = -b_lo / a_lo
    for k in range(n - 2, -1, -1):
        nxt = out[k + 1]
        out[k] = chi[k] if nxt > chi[k] else (clo[k] if nxt < clo[k] else nxt)
    return out


def tv1d(y: np.ndarray, lam: float) -> np.ndarray:
    """TotalVariation1DProx::ApplyVector, total_variation_1d.cc:12-24."""
    y = np.ascontiguousarray(y, dtype=np.float64)
    lib = _c_lib()
    if lib is None:
        return tv1d_py(y, float(lam))
    out = np.empty_like(y)
    rc = lib.oracle_tv1d(int(y.size), y.ctypes.data, float(lam), out.ctypes.data)
    assert rc == 0
    return out


# ---------------------------------------------------------------------------
# flattening (what affine::BuildAffineOperator produces, affine.cc:22-129)
# ---------------------------------------------------------------------------

def _term_matrix(t, rows: int):
    """The sparse matrix of one term's linear map (identity when map is None)."""
    if t.map is None:
        return sp.identity(rows, format="csr")
    return sp.csr_matrix(t.map.to_scipy())


def _affine_blocks(aff, var_cols: Dict[str, slice], ncols: int):
    """(matrix over the given column layout, constant vector) of one affine expression."""
    A = sp.csr_matrix((aff.rows, ncols))
    b = np.zeros(aff.rows)
    for t in aff.terms:
        if hasattr(t.operand, "var_id"):
            vid = t.operand.var_id
            if vid not in var_cols:
                continue
            sl = var_cols[vid]
            M = _term_matrix(t, aff.rows)
            E = sp.csr_matrix((np.ones(sl.stop - sl.start), (np.arange(sl.stop - sl.start),
                                                             np.arange(sl.start, sl.stop))),
                              shape=(sl.stop - sl.start, ncols))
            A = A + M @ E
        else:
            c = np.asarray(t.operand, dtype=np.float64).ravel()
            b += c if t.map is None else np.asarray(t.map.to_scipy() @ c).ravel()
    return sp.csr_matrix(A), b


def _diag_if_diagonal(M) -> Optional[np.ndarray]:
    M = sp.coo_matrix(M)
    if np.any(M.row != M.col):
        off = M.data[M.row != M.col]
        if np.any(off != 0):
            return None
    return np.asarray(sp.csr_matrix(M).diagonal()).ravel()


# ---------------------------------------------------------------------------
# operators
# ---------------------------------------------------------------------------

class _KktProx:
    """SumSquareProx (sum_square.cc:10-40) and AffineProx (affine.cc:19-49).

    Both assemble a symmetric block KKT matrix and eliminate it with BlockCholesky
    (block_cholesky.cc:119-137); the variables' part of chol_.Solve(b_ + v) is
        x = (2a H'H + A'A)^-1 (A'v - 2a H'g - c)
    (a = ProxFunction.alpha, entering as sqrt(2a) scaling of H and g, sum_square.cc:16,23,27;
    a = 0 for AFFINE / CONSTANT whose c = alpha * (1 x n map)', affine.cc:8-17,26-28).
    The elimination order NextKey picks only decides WHICH Schur complement is inverted
    (rows(H)^2 when rows(H) < n: the Woodbury side; else n^2) -- restated here by the same
    size test, with a dense Cholesky of that complement.
    """

    def __init__(self, f, H, g, A, matrix_free: bool = False):
        self.A = sp.csr_matrix(A)
        n = A.shape[1]
        a2 = 2.0 * f.alpha if f.type == "SUM_SQUARE" else 0.0
        self.c0 = np.zeros(n)
        if f.type == "SUM_SQUARE":
            self.c0 = -a2 * np.asarray(H.T @ g).ravel()
        elif f.type == "AFFINE":
            assert H.shape[0] == 1
            self.c0 = -f.alpha * np.asarray(H.todense()).ravel()
        self.matrix_free = bool(matrix_free)
        if self.matrix_free:
            # Same normal equations, solved without forming them: the reference densifies and inverts
            # the n x n Schur complement (sparse_matrix_impl.cc:60-78), which at n = 1e5 is 80 GB and
            # 3e14 flop.  Plain conjugate gradients on (2a H'H + A'A), warm-started from the previous
            # solution, run to a relative residual of 1e-13 -- below the 1e-12 of the device path it
            # checks, and far below the 0.3 % margin of the stop test.
            self.woodbury = False
            self.a2 = a2
            self.H = sp.csr_matrix(H) if a2 != 0 else None
            self.Ht = sp.csr_matrix(self.H.T) if a2 != 0 else None
            self.At = sp.csr_matrix(self.A.T)
            self.x_prev = np.zeros(n)
            self.cg_iterations = 0
            return
        AtA = (self.A.T @ self.A).tocsc()
        d = _diag_if_diagonal(AtA)
        rows_h = H.shape[0] if a2 != 0 else 0
        self.woodbury = bool(a2 != 0 and d is not None and np.all(d > 0) and 0 < rows_h < n)
        if self.woodbury:
            # (D + 2a H'H)^-1 = D^-1 - D^-1 H' (I/(2a) + H D^-1 H')^-1 H D^-1
            self.dinv = 1.0 / d
            self.H = sp.csr_matrix(H)
            Hd = self.H.toarray() if rows_h * n <= 5e7 else None
            if Hd is not None:
                S = (Hd * self.dinv) @ Hd.T
            else:
                S = (self.H @ sp.diags(self.dinv) @ self.H.T).toarray()
            S[np.diag_indices_from(S)] += 1.0 / a2
            self.chol = sla.cho_factor(S, lower=True)
        else:
            Msp = AtA + (a2 * (sp.csr_matrix(H).T @ sp.csr_matrix(H)) if a2 != 0 else 0 * AtA)
            self.mdiag = _diag_if_diagonal(Msp)      # every block scalar/diagonal: elementwise
            if self.mdiag is None:
                self.chol = sla.cho_factor(Msp.toarray(), lower=True)

    def _matvec(self, p: np.ndarray) -> np.ndarray:
        w = np.asarray(self.At @ np.asarray(self.A @ p).ravel()).ravel()
        if self.a2 != 0:
            w += self.a2 * np.asarray(self.Ht @ np.asarray(self.H @ p).ravel()).ravel()
        return w

    def _cg(self, rhs: np.ndarray, tol: float = 1e-13, max_iter: int = 2000) -> np.ndarray:
        x = self.x_prev.copy()
        r = rhs - self._matvec(x)
        p = r.copy()
        rr = float(r @ r)
        thresh = tol * float(np.sqrt(rhs @ rhs))
        k = 0
        while np.sqrt(rr) > thresh and k < max_iter:
            w = self._matvec(p)
            a = rr / float(p @ w)
            x += a * p
            r -= a * w
            rr_new = float(r @ r)
            p = r + (rr_new / rr) * p
            rr = rr_new
            k += 1
        assert np.sqrt(rr) <= thresh, "oracle CG did not converge"
        self.cg_iterations += k
        self.x_prev = x.copy()
        return x

    def apply(self, v: np.ndarray) -> np.ndarray:
        rhs = np.asarray(self.A.T @ v).ravel() + self.c0
        if self.matrix_free:
            return self._cg(rhs)
        if not self.woodbury:
            return rhs / self.mdiag if self.mdiag is not None else sla.cho_solve(self.chol, rhs)
        t = self.dinv * rhs
        z = sla.cho_solve(self.chol, np.asarray(self.H @ t).ravel())
        return t - self.dinv * np.asarray(self.H.T @ z).ravel()


class _VectorProx:
    """VectorProx::{Init,Apply} (vector_prox.cc:51-183): reduce
    argmin_x f(Hx + g) + 1/2||Ax - v||^2 to a plain prox when H'H and H A'A H' are scalar
    (InitScalar :51-70) or diagonal (InitDiagonal :72-116)."""

    def __init__(self, f, H, g, A):
        self.f = f
        H, A = sp.csr_matrix(H), sp.csr_matrix(A)
        n = H.shape[0]
        beta = _diag_if_diagonal(H.T @ H)
        gamma = _diag_if_diagonal(H @ A.T @ A @ H.T)
        assert beta is not None and gamma is not None, "Affine transformation is not scalar or diagonal"
        scalar = np.all(beta == beta[0]) and np.all(gamma == gamma[0]) and gamma[0] != 0
        self.g = g
        self.D = None
        if scalar:                                                     # :59-67
            b0, g0 = float(beta[0]), float(gamma[0])
            self.B = ((b0 / g0) * H) @ A.T
            self.C = (1 / b0) * H.T
            self.lam = np.full(n, f.alpha * b0 * b0 / g0)
            assert self.lam[0] >= 0
        else:                                                          # :84-115
            assert H.shape[0] == H.shape[1], "diagonal reduction restated for square H only"
            beta, gamma = beta.copy(), gamma.copy()
            lam, delta = np.zeros(n), np.zeros(n)
            for i in range(n):
                if gamma[i]:
                    lam[i] = f.alpha * beta[i] * beta[i] / gamma[i]
                else:
                    beta[i], gamma[i], delta[i] = 1.0, 1.0, 1.0
            self.B = H @ sp.diags(beta / gamma) @ A.T
            self.C = sp.diags(1.0 / beta) @ H.T
            AtA = _diag_if_diagonal(A.T @ A)
            assert AtA is not None
            inv = np.where(AtA != 0, 1.0 / np.where(AtA != 0, AtA, 1.0), 0.0)   # zero-safe Inverse
            self.D = sp.diags(inv) @ sp.diags(delta) @ A.T
            self.lam = lam
        self.scalar = scalar

    def _apply_epigraph(self, v: np.ndarray) -> np.ndarray:
        """Epigraph operators: arguments [x (n) | t (1)], projection onto {f(x) <= t}."""
        t = self.f.type
        n = self.f.args[0].rows
        if t in ORTHO_INVARIANT:
            m, k = self.f.mat_dims
            xo, to = ortho_invariant(t, True, m, k, 1.0, v[:n], float(v[n]))
            return np.concatenate([xo, [to]])
        if t == "EXP":                                   # elementwise: arguments [x (n) | t (n)]
            assert len(self.f.args) == 2 and self.f.args[1].rows == n
            xo, to = exp_epigraph(v[:n], v[n:])
            return np.concatenate([xo, to])
        if t == "SUM_KL_DIV":                            # arguments [x (n) | y (n) | t (1)]
            assert len(self.f.args) == 3 and self.f.args[1].rows == n and self.f.args[2].rows == 1
            xo, yo, to = kl_div_epigraph(v[:n], v[n:2 * n], float(v[2 * n]))
            return np.concatenate([xo, yo, [to]])
        assert len(self.f.args) == 2 and self.f.args[1].rows == 1
        x, s = v[:n], float(v[n])
        if t == "SUM_NEG_LOG":
            xo, to = neg_log_epigraph(x, s)
        elif t == "SUM_NEG_ENTR":
            xo, to = implicit_newton_epigraph(x, s, _neg_entr_eval, *SMOOTH[t])
        elif t == "LOG_SUM_EXP":
            xo, to = log_sum_exp_epigraph(x, s)
        elif t == "SUM_LARGEST":
            k = int(self.f.sum_largest_k)
            xo, to = bisection_epigraph(x, s, lambda z, lam: sum_largest_prox(z, lam, k),
                                        lambda z: _sum_largest_eval(z, k))
        elif t == "SUM_SQUARE":
            xo, to = sum_square_epigraph(x, s)
        elif t == "MAX":
            xo, to = max_epigraph(x, s)
        elif t in ZONE_TYPES:
            al, be, M = zone_params(self.f, n)
            assert np.ndim(al) == 0 or np.all(al == al[0]), "epigraph with vector alpha/beta"
            a0 = float(np.ravel(al)[0]); b0 = float(np.ravel(be)[0])
            xo, to = scaled_zone_epigraph(x, s, a0, b0, M)
        elif t in SMOOTH_EVAL:
            if t == "SUM_LOGISTIC":
                gradf, hessf, proj = _logistic_grad, _logistic_hess, (lambda z: z)
            else:
                gradf, hessf, proj = SMOOTH[t]
            xo, to = newton_epigraph(x, s, SMOOTH_EVAL[t], gradf, hessf, proj)
        else:
            raise ValueError("no restatement for epigraph type %s" % t)
        return np.concatenate([xo, [to]])

    def _apply_vector(self, v: np.ndarray) -> np.ndarray:
        t = self.f.type
        if self.f.epigraph:
            return self._apply_epigraph(v)
        if t in ZONE_TYPES:
            al, be, M = zone_params(self.f, v.size)
            return scaled_zone(v, self.lam, al, be, M)
        if t == "NON_NEGATIVE":
            return np.maximum(v, 0.0)                                   # non_negative.cc:5-9
        if t == "NORM_2":
            assert self.scalar
            if getattr(self.f, "axis", None) is not None:
                # has_axis (vector_prox.cc:150-177): ApplyVector once per column (axis 0) / row (axis 1) of
                # mat(arg 0, m, n), column-major (ToMatrix, vector_util.cc:202-209)
                m, n = self.f.axis_dims
                V = v.reshape((m, n), order="F")
                X = np.empty_like(V)
                if self.f.axis == 0:
                    for c in range(n):
                        X[:, c] = norm2_block(V[:, c], float(self.lam[0]))
                else:
                    for r in range(m):
                        X[r, :] = norm2_block(np.ascontiguousarray(V[r, :]), float(self.lam[0]))
                return X.ravel(order="F")
            return norm2_block(v, float(self.lam[0]))
        if t == "SUM_LOGISTIC":
            return newton_logistic(v, self.lam)
        if t in SMOOTH:
            return newton_prox(v, self.lam, *SMOOTH[t])
        if t == "SUM_NEG_LOG":
            return neg_log_prox(v, self.lam)
        if t == "TOTAL_VARIATION_1D":
            assert self.scalar
            return tv1d(v, float(self.lam[0]))
        sizes = [a.rows for a in self.f.args]
        offs = np.concatenate([[0], np.cumsum(sizes)])
        arg = lambda k: v[offs[k]:offs[k + 1]]
        if t in ORTHO_INVARIANT:
            m, k = self.f.mat_dims
            return ortho_invariant(t, False, m, k, float(self.lam[0]), v)[0]
        if t == "MAX":
            return max_prox(v, float(self.lam[0]))
        if t == "LOG_SUM_EXP":
            return log_sum_exp_prox(v, float(self.lam[0]))
        if t == "SUM_LARGEST":
            return sum_largest_prox(v, float(self.lam[0]), int(self.f.sum_largest_k))
        if t == "SUM_KL_DIV":
            x, y = kl_div_prox(self.lam[:sizes[0]], arg(0), arg(1))
            return np.concatenate([x, y])
        raise ValueError("no restatement for prox type %s" % t)

    def apply(self, v: np.ndarray) -> np.ndarray:
        vin = np.asarray(self.B @ v).ravel() + self.g                   # PreProcessInput :139-141
        x = self._apply_vector(vin)
        out = np.asarray(self.C @ (x - self.g)).ravel()                 # PostProcessOutput :143-145
        if self.D is not None:
            out = out + np.asarray(self.D @ v).ravel()
        return out


class _SocProx:
    """SecondOrderConeProx, second_order_cone.cc:30-110: rows of [t X] projected onto
    ||ax x_i + bx_i|| <= at t_i + bt_i; arguments (t, X), X in R^{m x n} column-major."""

    def __init__(self, f, H, g, A, arg_cols):
        m, n = f.soc_dims
        H, A = sp.csr_matrix(H), sp.csr_matrix(A)
        self.m, self.n = m, n
        # H is [at I on t ; ax I on x] up to the column order of the variables
        Hd = H.tocoo()
        self.tcols = np.empty(m, dtype=np.int64); self.xcols = np.empty(m * n, dtype=np.int64)
        at = ax = None
        for r, c, val in zip(Hd.row, Hd.col, Hd.data):
            if r < m:
                self.tcols[r] = c; at = val
            else:
                self.xcols[r - m] = c; ax = val
        ata = _diag_if_diagonal(A.T @ A)
        assert ata is not None and np.all(ata == ata[0]), "A'A not scalar matrix"
        self.AT = (1.0 / ata[0]) * A.T                                   # InitConstraints :93-104
        g = np.zeros(m + m * n) if g is None or np.size(g) == 0 else np.asarray(g, dtype=np.float64)
        self.a = at / abs(ax)                                           # InitArgs :76-90
        self.bx = g[m:] / ax
        self.bt = g[:m] / abs(ax)
        self.ncols = H.shape[1]

    def apply(self, v: np.ndarray) -> np.ndarray:
        u = np.asarray(self.AT @ v).ravel()
        X = (u[self.xcols] + self.bx).reshape((self.m, self.n), order="F")
        t = u[self.tcols] + self.bt / self.a
        beta = self.a
        with np.errstate(all="ignore"):
            vn = np.sqrt((X * X).sum(axis=1))
            alpha = (1 / (beta * beta + 1)) * (beta * beta + beta * t / vn)   # ApplyProjection :54-70
        for i in range(self.m):
            if np.isnan(alpha[i]) or alpha[i] > 1:
                alpha[i] = 1
            elif alpha[i] < 0:
                alpha[i] = 0
                t[i] = 0
            else:
                t[i] = (1 / beta) * alpha[i] * vn[i]
        X = alpha[:, None] * X
        out = np.zeros(self.ncols)
        out[self.xcols] = X.ravel(order="F") - self.bx
        out[self.tcols] = t - self.bt / self.a
        return out


def _make_operator(f, H, g, A, matrix_free: bool = False):
    if f.type == "SECOND_ORDER_CONE":
        return _SocProx(f, H, g, A, None)
    if f.type in KKT_TYPES and not f.epigraph:
        return _KktProx(f, H, g, A, matrix_free)
    return _VectorProx(f, H, g, A)


# ---------------------------------------------------------------------------
# ProxADMMSolver
# ---------------------------------------------------------------------------

class ProxADMM:
    """ProxADMMSolver (prox_admm.cc).  `problem` is an epsilon_b200.problem.ProxAffineProblem
    (the flattened prox-affine form both sides of the boundary share)."""

    def __init__(self, problem, params: Optional[dict] = None):
        self.problem = problem
        self.params = dict(DEFAULT_PARAMS)
        self.params.update(params or {})
        assert self.params["rho"] == 1.0, "CHECK_EQ(1, rho), prox_admm.cc:50"
        self._init()

    # InitConstraints (:25-43) + InitProxOperators (:45-94) + InitVariables (:96-108)
    def _init(self):
        p = self.problem
        sizes = p.variables()
        self.N = len(p.prox)
        self.prox_vars: List[List[str]] = [p.prox_variables(i) for i in range(self.N)]
        owner = {}
        for i, vs in enumerate(self.prox_vars):
            for v in vs:
                assert v not in owner, "variable %s in two prox functions" % v
                owner[v] = i
        coff = np.concatenate([[0], np.cumsum([c.rows for c in p.constraints])]).astype(int)
        self.m = int(coff[-1])                                         # A_.m()
        in_constraint = set()
        for c in p.constraints:
            for t in c.terms:
                if hasattr(t.operand, "var_id"):
                    in_constraint.add(t.operand.var_id)
        self.n = int(sum(sizes[v] for v in in_constraint))             # A_.n()
        self.b = np.zeros(self.m)
        self.Ai, self.ops, self.cols = [], [], []
        for i, f in enumerate(p.prox):
            cols, off = {}, 0
            for v in self.prox_vars[i]:
                cols[v] = slice(off, off + sizes[v])
                off += sizes[v]
            self.cols.append(cols)
            # H, g: the arguments stacked (keys arg:0, arg:1, ...)
            Hs, gs = [], []
            for a in f.args:
                Ha, ga = _affine_blocks(a, cols, off)
                Hs.append(Ha)
                gs.append(ga)
            H = sp.vstack(Hs, format="csr") if Hs else sp.csr_matrix((0, off))
            g = np.concatenate(gs) if gs else np.zeros(0)
            # A_i: columns of A_ that belong to f_i's variables, times sqrt(rho) (:68-79)
            in_c = {v: s for v, s in cols.items() if v in in_constraint}
            rows = []
            for c in p.constraints:
                Ac, _ = _affine_blocks(c, in_c, off)
                rows.append(Ac)
            A = sp.vstack(rows, format="csr") if rows else sp.csr_matrix((0, off))
            self.Ai.append(A)
            # "kkt": "cg" is an oracle-only switch (not a SolverParams field): matrix-free KKT solves
            # for shapes whose n x n factor the reference could not form
            self.ops.append(_make_operator(f, H, g, A, self.params.get("kkt") == "cg"))
        for k, c in enumerate(p.constraints):
            _, bc = _affine_blocks(c, {}, 0)
            self.b[coff[k]:coff[k + 1]] = bc
        self.reset()

    def reset(self):
        self.u = np.zeros(self.m)
        self.x = [np.zeros(A.shape[1]) for A in self.Ai]
        for op in self.ops:
            if getattr(op, "matrix_free", False):
                op.x_prev[:] = 0.0
        self.y = [np.zeros(self.m) for _ in range(self.N)]             # empty BlockVector = 0
        self.y_prev = [np.zeros(self.m) for _ in range(self.N)]
        self.iter = 0
        self.status = {"state": 0, "num_iterations": 0}

    def iteration(self):
        """One pass of prox_admm.cc:135-147."""
        self.y_prev = [yi.copy() for yi in self.y]
        self.u = self.u - self.b
        for i in range(self.N):
            self.u = self.u - self.y[i]
        for i in range(self.N):
            self.u = self.u + self.y[i]
            self.x[i] = self.ops[i].apply(self.u)
            self.y[i] = np.asarray(self.Ai[i] @ self.x[i]).ravel()
            self.u = self.u - self.y[i]

    def compute_residuals(self) -> dict:
        """prox_admm.cc:178-217."""
        abs_tol, rel_tol, rho = self.params["abs_tol"], self.params["rel_tol"], self.params["rho"]
        ax_b = self.b.copy()
        max_ax = float(np.linalg.norm(self.b))
        for i in range(self.N):
            ai_xi = np.asarray(self.Ai[i] @ self.x[i]).ravel()
            max_ax = max(max_ax, float(np.linalg.norm(ai_xi)))
            ax_b += ai_xi
        s2 = 0.0
        diff = np.zeros(self.m)
        for i in range(self.N - 2, -1, -1):
            diff = diff + (self.y[i + 1] - self.y_prev[i + 1])
            si = float(np.linalg.norm(self.Ai[i].T @ diff))
            s2 += si * si
        atu2 = sum(float(np.sum(np.asarray(A.T @ self.u) ** 2)) for A in self.Ai)
        st = {
            "r_norm": float(np.linalg.norm(ax_b)),
            "s_norm": rho * float(np.sqrt(s2)),
            "epsilon_primal": abs_tol * np.sqrt(self.m) + rel_tol * max_ax,
            "epsilon_dual": abs_tol * np.sqrt(self.n) + rel_tol * rho * float(np.sqrt(atu2)),
        }
        ok = st["r_norm"] <= st["epsilon_primal"] and st["s_norm"] <= st["epsilon_dual"]
        st["state"] = STATE_OPTIMAL if ok else STATE_RUNNING
        st["num_iterations"] = self.iter
        self.status = st
        return st

    def solution(self) -> Dict[str, np.ndarray]:
        """GetSolution, prox_admm.cc:171-176."""
        out = {}
        for i in range(self.N):
            for v, sl in self.cols[i].items():
                out[v] = self.x[i][sl].copy()
        return out

    def solve(self) -> dict:
        """ProxADMMSolver::Solve, prox_admm.cc:131-169.  Returns status, per-epoch trace, x."""
        self.reset()
        trace = []
        max_it, epoch = int(self.params["max_iterations"]), int(self.params["epoch_iterations"])
        self.iter = 0
        while self.iter < max_it:
            self.iteration()
            if self.iter % epoch == 0:
                trace.append(dict(self.compute_residuals()))
                if self.status["state"] == STATE_OPTIMAL:
                    break
            self.iter += 1
        if self.iter == max_it:
            self.compute_residuals()
            self.status["state"] = STATE_MAX_ITERATIONS
            trace.append(dict(self.status))
        return {"status": dict(self.status), "trace": trace, "x": self.solution(),
                "m": self.m, "n": self.n, "N": self.N}


def eval_prox(f, lam: float, v: Dict[str, np.ndarray]) -> Dict[str, np.ndarray]:
    """EvalProx, python/epopt/solvemodule.cc:189-242: A = (1/sqrt(lam)) I per variable,
    v <- A v, one Apply."""
    from epsilon_b200.problem import Affine, LinearMapSpec, ProxAffineProblem, Term, Variable
    prob = ProxAffineProblem([f], [])
    sizes = prob.variables()
    a = 1.0 / np.sqrt(lam)
    order = sorted(sizes)
    cons = [Affine(sizes[k], [Term(Variable(k, sizes[k]), LinearMapSpec.Scalar(a, sizes[k]))])
            for k in order]
    s = ProxADMM(ProxAffineProblem([f], cons))
    u = np.concatenate([a * np.asarray(v[k], dtype=np.float64).ravel() for k in order])
    x = s.ops[0].apply(u)
    return {k: x[sl].copy() for k, sl in s.cols[0].items()}


# ---------------------------------------------------------------------------
# ProxADMMTwoBlockSolver (src/epsilon/algorithms/prox_admm_two_block.cc)
# ---------------------------------------------------------------------------

class ProxADMMTwoBlock:
    """Consensus form: x_i = prox_{f_i}(z - u) for all i, z = projection of x + u onto
    {H z + b = 0} (ZeroProx, prox/zero.cc:10-36), u += x - z  (prox_admm_two_block.cc:96-133).
    Every prox sees A = sqrt(rho) * I on its own variables (:66-71)."""

    def __init__(self, problem, params: Optional[dict] = None):
        self.problem = problem
        self.params = dict(DEFAULT_PARAMS)
        self.params.update(params or {})
        assert self.params["rho"] == 1.0
        p = problem
        sizes = p.variables()
        self.N = len(p.prox)
        # flat layout: variables grouped by owning prox function, ids sorted inside a group
        self.cols: Dict[str, slice] = {}
        off = 0
        self.prox_cols = []
        for i in range(self.N):
            loc = {}
            start = off
            for v in p.prox_variables(i):
                self.cols[v] = slice(off, off + sizes[v])
                loc[v] = slice(off - start, off - start + sizes[v])
                off += sizes[v]
            self.prox_cols.append((start, off, loc))
        self.n_all = off
        # InitConstraints (:21-49): H over the constraint variables, z = 0 for them
        rows, bs = [], []
        in_constraint = set()
        for c in p.constraints:
            Hc, bc = _affine_blocks(c, self.cols, self.n_all)
            rows.append(Hc)
            bs.append(bc)
            for t in c.terms:
                if hasattr(t.operand, "var_id"):
                    in_constraint.add(t.operand.var_id)
        self.H = sp.vstack(rows, format="csr") if rows else sp.csr_matrix((0, self.n_all))
        self.hb = np.concatenate(bs) if bs else np.zeros(0)
        self.m = self.H.shape[0]
        self.n = int(sum(sizes[v] for v in in_constraint))
        self.mask = np.zeros(self.n_all)
        for v in in_constraint:
            self.mask[self.cols[v]] = 1.0
        if self.m:
            HHt = (self.H @ self.H.T).tocsc()
            d = _diag_if_diagonal(HHt)
            self.hh_diag = d
            self.hh_chol = None if d is not None else sla.cho_factor(HHt.toarray(), lower=True)
        # InitProxOperators (:51-84)
        self.ops = []
        for i, f in enumerate(p.prox):
            start, end, loc = self.prox_cols[i]
            Hs, gs = [], []
            for a in f.args:
                Ha, ga = _affine_blocks(a, loc, end - start)
                Hs.append(Ha)
                gs.append(ga)
            H = sp.vstack(Hs, format="csr") if Hs else sp.csr_matrix((0, end - start))
            g = np.concatenate(gs) if gs else np.zeros(0)
            self.ops.append(_make_operator(f, H, g, sp.identity(end - start, format="csr")))
        self.reset()

    def reset(self):
        self.z = np.zeros(self.n_all)
        self.z_prev = np.zeros(self.n_all)
        self.u = np.zeros(self.n_all)
        self.x = np.zeros(self.n_all)
        self.iter = 0
        self.status = {"state": 0, "num_iterations": 0}

    def _project(self, v):
        """ZeroProx::Apply with A = I: v - H'(HH')^-1 (Hv + b), restricted to H's columns."""
        if self.m == 0:
            return v * self.mask
        r = np.asarray(self.H @ v).ravel() + self.hb
        y = r / self.hh_diag if self.hh_diag is not None else sla.cho_solve(self.hh_chol, r)
        return (v - np.asarray(self.H.T @ y).ravel()) * self.mask

    def iteration(self):
        self.z_prev = self.z.copy()
        zu = self.z - self.u
        x = np.zeros(self.n_all)
        for i in range(self.N):
            start, end, _ = self.prox_cols[i]
            x[start:end] = self.ops[i].apply(zu[start:end])
        self.x = x
        self.z = self._project(self.x + self.u)
        self.u = self.u + (self.x - self.z)

    def compute_residuals(self):
        """prox_admm_two_block.cc:135-156 (both tolerances use sqrt(n_))."""
        abs_tol, rel_tol, rho = self.params["abs_tol"], self.params["rel_tol"], self.params["rho"]
        nrm = lambda a: float(np.sqrt(np.sum(a * a)))
        st = {"r_norm": nrm(self.x - self.z), "s_norm": rho * nrm(self.z - self.z_prev),
              "epsilon_primal": abs_tol * np.sqrt(self.n) + rel_tol * max(nrm(self.x), nrm(self.z)),
              "epsilon_dual": abs_tol * np.sqrt(self.n) + rel_tol * rho * nrm(self.u)}
        ok = st["r_norm"] <= st["epsilon_primal"] and st["s_norm"] <= st["epsilon_dual"]
        st["state"] = STATE_OPTIMAL if ok else STATE_RUNNING
        st["num_iterations"] = self.iter
        self.status = st
        return st

    def solution(self):
        return {v: self.x[sl].copy() for v, sl in self.cols.items()}

    def solve(self):
        self.reset()
        trace = []
        max_it, epoch = int(self.params["max_iterations"]), int(self.params["epoch_iterations"])
        self.iter = 0
        while self.iter < max_it:
            self.iteration()
            if self.iter % epoch == 0:
                trace.append(dict(self.compute_residuals()))
                if self.status["state"] == STATE_OPTIMAL:
                    break
            self.iter += 1
        if self.iter == max_it:
            self.compute_residuals()
            self.status["state"] = STATE_MAX_ITERATIONS
            trace.append(dict(self.status))
        return {"status": dict(self.status), "trace": trace, "x": self.solution(),
                "m": self.m, "n": self.n, "N": self.N}
