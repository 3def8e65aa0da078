"""Prox-affine problem description: the data model both sides of the boundary share.

A problem in prox-affine form is  ``minimize sum_i f_i(H_i x + g_i)  s.t.  sum_j A_j x_j + b = 0``
(reference: docs/solver.rst:11-106).  In the reference this is the protobuf
``Problem`` of proto/epsilon/expression.proto:272-281 whose expression trees are
flattened at Init by ``affine::BuildAffineOperator`` (src/epsilon/affine/affine.cc:22-129)
into block matrices keyed by ``constraint:i`` / ``arg:i`` rows and variable-id columns.

This module holds that *flattened* form as plain Python objects and can
  * write it as a protobuf-text ``Problem`` plus raw constant bytes in the wire
    layout of python/epopt/constant.py:10-41 (dense = column-major f64; sparse =
    int32 colptr | int32 rowidx | f64 values), which the compiled reference
    (oracle/_ref/epsilon_ref) consumes, and
  * be handed to the device solver through the C-ABI (epsilon_b200/solver.py).

Only the expression node types the affine builder accepts are produced:
ADD, CONSTANT, LINEAR_MAP, VARIABLE (affine.cc:102-108).
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Union

import numpy as np
import scipy.sparse as sp

# ProxFunction.Type values, proto/epsilon/expression.proto:123-170.
PROX_TYPES = {
    "AFFINE": 1,
    "CONSTANT": 2,
    "ZERO": 10,
    "SUM_SQUARE": 11,
    "NON_NEGATIVE": 20,
    "NORM_1": 21,
    "SUM_DEADZONE": 22,
    "SUM_HINGE": 24,
    "SUM_LOGISTIC": 27,
    "SUM_QUANTILE": 31,
    "NORM_2": 102,
    "TOTAL_VARIATION_1D": 106,
}

# linear_map::ImplType order, src/epsilon/linear/linear_map.h (DENSE < SPARSE < DIAGONAL < SCALAR < KRONECKER).
MAP_DENSE, MAP_SPARSE, MAP_DIAGONAL, MAP_SCALAR, MAP_KRONECKER = range(5)


@dataclass
class LinearMapSpec:
    """One atomic linear map (proto LinearMap, expression.proto:96-121)."""

    kind: int
    m: int
    n: int
    dense: Optional[np.ndarray] = None  # (m, n) float64, any order (written column-major)
    sparse: Optional[sp.csc_matrix] = None
    diag: Optional[np.ndarray] = None
    scalar: float = 1.0
    kron: Optional[Sequence["LinearMapSpec"]] = None  # (A, B) for A (x) B
    trans: bool = False  # DENSE only: stored data is the transpose's (shares storage)
    name: str = ""  # data_location stem for the reference

    @staticmethod
    def Dense(a: np.ndarray, name: str = "") -> "LinearMapSpec":
        a = np.asarray(a, dtype=np.float64)
        return LinearMapSpec(MAP_DENSE, a.shape[0], a.shape[1], dense=a, name=name)

    @staticmethod
    def Sparse(a, name: str = "") -> "LinearMapSpec":
        a = sp.csc_matrix(a, dtype=np.float64)
        a.sort_indices()
        return LinearMapSpec(MAP_SPARSE, a.shape[0], a.shape[1], sparse=a, name=name)

    @staticmethod
    def Diagonal(d: np.ndarray, name: str = "") -> "LinearMapSpec":
        d = np.asarray(d, dtype=np.float64).ravel()
        return LinearMapSpec(MAP_DIAGONAL, d.size, d.size, diag=d, name=name)

    @staticmethod
    def Scalar(alpha: float, n: int) -> "LinearMapSpec":
        return LinearMapSpec(MAP_SCALAR, n, n, scalar=float(alpha))

    @staticmethod
    def Identity(n: int) -> "LinearMapSpec":
        return LinearMapSpec.Scalar(1.0, n)

    @staticmethod
    def Kronecker(a: "LinearMapSpec", b: "LinearMapSpec") -> "LinearMapSpec":
        return LinearMapSpec(MAP_KRONECKER, a.m * b.m, a.n * b.n, kron=(a, b))

    def to_scipy(self):
        """The map as a scipy/numpy operator (used by the numpy oracle and tests)."""
        if self.kind == MAP_DENSE:
            return self.dense
        if self.kind == MAP_SPARSE:
            return self.sparse
        if self.kind == MAP_DIAGONAL:
            return sp.diags([self.diag], [0], format="csc")
        if self.kind == MAP_SCALAR:
            return self.scalar * sp.identity(self.n, format="csc")
        a, b = self.kron
        return sp.kron(sp.csc_matrix(a.to_scipy()), sp.csc_matrix(b.to_scipy()), format="csc")

    def apply(self, x: np.ndarray) -> np.ndarray:
        return np.asarray(self.to_scipy() @ x).ravel()


@dataclass
class Variable:
    var_id: str
    size: int


@dataclass
class Term:
    """``map * operand`` where operand is a Variable or a constant vector.

    map=None means the identity (a bare VARIABLE / CONSTANT node).
    """

    operand: Union[Variable, np.ndarray]
    map: Optional[LinearMapSpec] = None
    name: str = ""  # data_location stem when operand is a constant


@dataclass
class Affine:
    """Sum of terms = one ADD node; ``rows`` is its dimension."""

    rows: int
    terms: List[Term] = field(default_factory=list)


@dataclass
class ProxFunction:
    """One objective term f_i (proto ProxFunction, expression.proto:123-196)."""

    type: str
    args: List[Affine]
    alpha: float = 1.0
    epigraph: bool = False
    # scaled_zone_params.m for SUM_DEADZONE (scaled_zone.cc:52-56)
    zone_m: float = 0.0
    # SUM_QUANTILE alpha/beta vectors (scaled_zone.cc:62-71)
    quantile_alpha: Optional[np.ndarray] = None
    quantile_beta: Optional[np.ndarray] = None


@dataclass
class ProxAffineProblem:
    prox: List[ProxFunction]
    constraints: List[Affine]
    name: str = ""

    def variables(self) -> Dict[str, int]:
        out: Dict[str, int] = {}
        for aff in [a for f in self.prox for a in f.args] + list(self.constraints):
            for t in aff.terms:
                if isinstance(t.operand, Variable):
                    out[t.operand.var_id] = t.operand.size
        return out

    def prox_variables(self, i: int) -> List[str]:
        """Variables of f_i in variable-id order (VariableSet is a std::set ordered by
        VariableIdCompare, expression_util.cc:5-9,20-31)."""
        seen = set()
        for aff in self.prox[i].args:
            for t in aff.terms:
                if isinstance(t.operand, Variable):
                    seen.add(t.operand.var_id)
        return sorted(seen)


# ---------------------------------------------------------------------------
# Protobuf text writer (for the compiled reference only)
# ---------------------------------------------------------------------------


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

    def dense_constant(self, a: np.ndarray, stem: str) -> str:
        a = np.asarray(a, dtype=np.float64)
        if a.ndim == 1:
            a = a.reshape(-1, 1)
        loc = self.store(stem, a.tobytes(order="F"))
        return 'constant { constant_type: DENSE_MATRIX m: %d n: %d data_location: "%s" }' % (
            a.shape[0], a.shape[1], loc)

    def sparse_constant(self, a: sp.csc_matrix, stem: str) -> str:
        payload = (a.indptr.astype(np.int32).tobytes() + a.indices.astype(np.int32).tobytes()
                   + a.data.astype(np.float64).tobytes())
        loc = self.store(stem, payload)
        return ('constant { constant_type: SPARSE_MATRIX m: %d n: %d nnz: %d data_location: "%s" }'
                % (a.shape[0], a.shape[1], a.nnz, loc))

    def linear_map(self, lm: LinearMapSpec) -> str:
        if lm.kind == MAP_DENSE:
            return "linear_map_type: DENSE_MATRIX m: %d n: %d %s" % (
                lm.m, lm.n, self.dense_constant(lm.dense, lm.name or "dense"))
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
                self.size(c.size), self.dense_constant(c, t.name or "const"))
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
        for a in f.args:
            pf += " arg_size { dim: %d dim: 1 }" % a.rows
        if f.type == "SUM_DEADZONE":
            pf += " scaled_zone_params { m: %s }" % repr(float(f.zone_m))
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
                            params: Optional[dict] = None) -> None:
    """Write <out_dir>/problem.pbtxt, params.pbtxt and data/ for oracle/_ref/epsilon_ref."""
    os.makedirs(out_dir, exist_ok=True)
    w = _RefWriter(os.path.join(out_dir, "data"))
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
            f.write("%s: %s\n" % (k, str(v).lower() if isinstance(v, bool) else repr(v)))


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
