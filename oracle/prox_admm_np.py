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
