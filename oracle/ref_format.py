"""Write a ProxAffineProblem in the reference's own input format.

TEST INFRASTRUCTURE (used by oracle/ref_runner.py only).  Produces a protobuf-text
``Problem`` (proto/epsilon/expression.proto:272-281) plus raw constant bytes in the wire
layout of python/epopt/constant.py:10-41 (dense = column-major f64; sparse = int32 colptr |
int32 rowidx | f64 values), which the compiled reference (oracle/_ref/epsilon_ref) consumes.
Only the expression node types the affine builder accepts are produced: ADD, CONSTANT,
LINEAR_MAP, VARIABLE (src/epsilon/affine/affine.cc:102-108).
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import numpy as np
import scipy.sparse as sp

from epsilon_b200.problem import (MAP_DENSE, MAP_DIAGONAL, MAP_SCALAR, MAP_SPARSE, Affine,
                                  LinearMapSpec, ProxAffineProblem, ProxFunction, Term, Variable)


class _RefWriter:
    def __init__(self, data_dir: str):
        self.data_dir = data_dir
        self.count = 0
        os.makedirs(data_dir, exist_ok=True)

    def store(self, stem: str, payload: bytes) -> str:
        loc = "%s_%d" % (stem or "c", self.count)
        self.count += 1
        with open(os.path.join(self.data_dir, loc), "wb") as f:
            f.write(payload)
        return loc

    def size(self, m: int, n: int = 1) -> str:
        return "size { dim: %d dim: %d }" % (m, n)

    def dense_constant_body(self, a: np.ndarray, stem: str, parameter_id: str = "") -> str:
        a = np.asarray(a, dtype=np.float64)
        if a.ndim == 1:
            a = a.reshape(-1, 1)
        loc = self.store(stem, a.tobytes(order="F"))
        body = 'constant_type: DENSE_MATRIX m: %d n: %d data_location: "%s"' % (a.shape[0], a.shape[1], loc)
        if parameter_id:
            body += ' parameter_id: "%s"' % parameter_id
        return body

    def dense_constant(self, a: np.ndarray, stem: str, parameter_id: str = "") -> str:
        return "constant { %s }" % self.dense_constant_body(a, stem, parameter_id)

    def sparse_constant(self, a: sp.csc_matrix, stem: str) -> str:
        payload = (a.indptr.astype(np.int32).tobytes() + a.indices.astype(np.int32).tobytes()
                   + a.data.astype(np.float64).tobytes())
        loc = self.store(stem, payload)
        return ('constant { constant_type: SPARSE_MATRIX m: %d n: %d nnz: %d data_location: "%s" }'
                % (a.shape[0], a.shape[1], a.nnz, loc))

    def linear_map(self, lm: LinearMapSpec) -> str:
        if lm.kind == MAP_DENSE:
            return "linear_map_type: DENSE_MATRIX m: %d n: %d %s" % (
                lm.m, lm.n, self.dense_constant(lm.dense, lm.name or "dense", getattr(lm, "parameter_id", "")))
        if lm.kind == MAP_SPARSE:
            return "linear_map_type: SPARSE_MATRIX m: %d n: %d %s" % (
                lm.m, lm.n, self.sparse_constant(lm.sparse, lm.name or "sparse"))
        if lm.kind == MAP_DIAGONAL:
            # DiagonalMatrix factory reads a dense n x 1 constant (linear_map.cc:64-70)
            return "linear_map_type: DIAGONAL_MATRIX m: %d n: %d %s" % (
                lm.m, lm.n, self.dense_constant(lm.diag, lm.name or "diag"))
        if lm.kind == MAP_SCALAR:
            return "linear_map_type: SCALAR m: %d n: %d scalar: %s" % (lm.m, lm.n, repr(lm.scalar))
        a, b = lm.kron
        return "linear_map_type: KRONECKER_PRODUCT m: %d n: %d arg { %s } arg { %s }" % (
            lm.m, lm.n, self.linear_map(a), self.linear_map(b))

    def term(self, t: Term, rows: int) -> str:
        if isinstance(t.operand, Variable):
            leaf = 'expression_type: VARIABLE %s variable { variable_id: "%s" }' % (
                self.size(t.operand.size), t.operand.var_id)
        else:
            c = np.asarray(t.operand, dtype=np.float64).ravel()
            leaf = "expression_type: CONSTANT %s %s" % (
                self.size(c.size), self.dense_constant(c, t.name or "const", getattr(t, "parameter_id", "")))
        if t.map is None:
            return leaf
        return "expression_type: LINEAR_MAP %s linear_map { %s } arg { %s }" % (
            self.size(rows), self.linear_map(t.map), leaf)

    def affine(self, aff: Affine) -> str:
        if len(aff.terms) == 1:
            return self.term(aff.terms[0], aff.rows)
        body = " ".join("arg { %s }" % self.term(t, aff.rows) for t in aff.terms)
        return "expression_type: ADD %s %s" % (self.size(aff.rows), body)

    def prox_function(self, f: ProxFunction) -> str:
        pf = "prox_function_type: %s alpha: %s" % (f.type, repr(float(f.alpha)))
        if f.epigraph:
            pf += " epigraph: true"
        for k, a in enumerate(f.args):
            if f.type == "SECOND_ORDER_CONE" and k == 1:
                pf += " arg_size { dim: %d dim: %d }" % tuple(f.soc_dims)
            elif f.mat_dims is not None and k == 0:
                pf += " arg_size { dim: %d dim: %d }" % tuple(f.mat_dims)
            elif getattr(f, "axis", None) is not None and k == 0:
                pf += " arg_size { dim: %d dim: %d }" % tuple(f.axis_dims)
            else:
                pf += " arg_size { dim: %d dim: 1 }" % a.rows
        if getattr(f, "axis", None) is not None:
            pf += " has_axis: true axis: %d" % int(f.axis)
        if f.type == "SUM_DEADZONE":
            pf += " scaled_zone_params { m: %s }" % repr(float(f.zone_m))
        if f.type == "SUM_LARGEST":
            pf += " sum_largest_params { k: %d }" % int(f.sum_largest_k)
        if f.type == "SUM_QUANTILE":
            qa = "expression_type: CONSTANT %s %s" % (
                self.size(f.quantile_alpha.size), self.dense_constant(f.quantile_alpha, "qalpha"))
            qb = "expression_type: CONSTANT %s %s" % (
                self.size(f.quantile_beta.size), self.dense_constant(f.quantile_beta, "qbeta"))
            pf += " scaled_zone_params { alpha_expr { %s } beta_expr { %s } }" % (qa, qb)
        args = " ".join("arg { %s }" % self.affine(a) for a in f.args)
        return "expression_type: PROX_FUNCTION %s prox_function { %s } %s" % (
            self.size(1), pf, args)


def write_reference_problem(problem: ProxAffineProblem, out_dir: str,
                            params: Optional[dict] = None,
                            reparam: Optional[Dict[str, np.ndarray]] = None) -> None:
    """Write <out_dir>/problem.pbtxt, params.pbtxt and data/ for oracle/_ref/epsilon_ref.
    `reparam`: {parameter_id: new constant vector} applied by the harness' resolve modes through
    Solver::SetParameterValue between two Solve() calls."""
    os.makedirs(out_dir, exist_ok=True)
    w = _RefWriter(os.path.join(out_dir, "data"))
    if reparam:
        os.makedirs(os.path.join(out_dir, "reparam"), exist_ok=True)
        for pid, val in reparam.items():
            with open(os.path.join(out_dir, "reparam", pid), "w") as f:
                val = np.asarray(val, dtype=np.float64)
                f.write(w.dense_constant_body(val if val.ndim == 2 else val.ravel(), "param_" + pid, pid) + "\n")
    obj = "expression_type: ADD %s %s" % (
        w.size(1), " ".join("arg { %s }" % w.prox_function(f) for f in problem.prox))
    parts = ["objective { %s }" % obj]
    for c in problem.constraints:
        parts.append("constraint { expression_type: INDICATOR %s cone { cone_type: ZERO } arg { %s } }"
                     % (w.size(1), w.affine(c)))
    with open(os.path.join(out_dir, "problem.pbtxt"), "w") as f:
        f.write("\n".join(parts) + "\n")
    with open(os.path.join(out_dir, "params.pbtxt"), "w") as f:
        for k, v in (params or {}).items():
            if isinstance(v, bool):
                txt = str(v).lower()
            elif isinstance(v, str):
                txt = v            # enum identifier (e.g. solver: PROX_ADMM_TWO_BLOCK)
            else:
                txt = repr(v)
            f.write("%s: %s\n" % (k, txt))


def write_reference_prox(f: ProxFunction, lam: float, v: Dict[str, np.ndarray], out_dir: str) -> None:
    """Inputs of the harness' `prox` mode (EvalProx, solvemodule.cc:189-242)."""
    os.makedirs(os.path.join(out_dir, "v"), exist_ok=True)
    w = _RefWriter(os.path.join(out_dir, "data"))
    with open(os.path.join(out_dir, "f_expr.pbtxt"), "w") as fh:
        fh.write(w.prox_function(f) + "\n")
    with open(os.path.join(out_dir, "lambda.txt"), "w") as fh:
        fh.write(repr(float(lam)))
    for var_id, val in v.items():
        np.asarray(val, dtype=np.float64).tofile(os.path.join(out_dir, "v", var_id))
