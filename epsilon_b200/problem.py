"""Prox-affine problem description: the data model both sides of the boundary share.

A problem in prox-affine form is  ``minimize sum_i f_i(H_i x + g_i)  s.t.  sum_j A_j x_j + b = 0``
(reference: docs/solver.rst:11-106).  In the reference this is the protobuf
``Problem`` of proto/epsilon/expression.proto:272-281 whose expression trees are
flattened at Init by ``affine::BuildAffineOperator`` (src/epsilon/affine/affine.cc:22-129)
into block matrices keyed by ``constraint:i`` / ``arg:i`` rows and variable-id columns.

This module holds that *flattened* form as plain Python objects; it is handed to the
device solver through the C-ABI (epsilon_b200/solver.py).  (The test infrastructure can
also write it as a protobuf-text ``Problem`` for the compiled reference: oracle/ref_format.py.)

Only the expression node types the affine builder accepts are produced:
ADD, CONSTANT, LINEAR_MAP, VARIABLE (affine.cc:102-108).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Tuple, Dict, List, Optional, Sequence, Union

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
    "SUM_EXP": 23,
    "SUM_HINGE": 24,
    "SUM_INV_POS": 25,
    "SUM_KL_DIV": 26,
    "SUM_LOGISTIC": 27,
    "SUM_NEG_ENTR": 28,
    "SUM_NEG_LOG": 29,
    "SUM_QUANTILE": 31,
    "EXP": 32,
    "LOG_SUM_EXP": 100,
    "MAX": 101,
    "NORM_2": 102,
    "SECOND_ORDER_CONE": 104,
    "SUM_LARGEST": 105,
    "TOTAL_VARIATION_1D": 106,
    "LAMBDA_MAX": 200,
    "NEG_LOG_DET": 202,
    "NORM_NUCLEAR": 203,
    "SEMIDEFINITE": 204,
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
    parameter_id: str = ""  # DENSE only: Constant.parameter_id of the matrix data (a cvxpy Parameter)
    # DENSE only: the matrix already lives on the device, column-major f64; (pointer, owner object kept alive)
    device_ptr: Optional[int] = None
    device_owner: object = None

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
    # Constant.parameter_id (proto/epsilon/expression.proto:22): a cvxpy Parameter, replaceable on a live
    # solver through Solver::SetParameterValue (solver.cc:109-116) / eps_solver_set_constant
    parameter_id: str = ""


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
    # sum_largest_params.k for SUM_LARGEST (sum_largest.cc:12)
    sum_largest_k: int = 0
    # SECOND_ORDER_CONE: (m, n) of the matrix argument X (arg_size(1), second_order_cone.cc:31-35)
    soc_dims: Optional[Tuple[int, int]] = None
    # matrix functions (LAMBDA_MAX, NEG_LOG_DET, NORM_NUCLEAR, SEMIDEFINITE): (m, n) of the matrix argument
    # mat(x, m, n), column-major (arg_size(0), ortho_invariant.cc:12-16)
    mat_dims: Optional[Tuple[int, int]] = None
    # has_axis / axis (expression.proto:193-195): apply to every column (0) or row (1) of mat(arg0, *axis_dims)
    axis: Optional[int] = None
    axis_dims: Optional[Tuple[int, int]] = None


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
