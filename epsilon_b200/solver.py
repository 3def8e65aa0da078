"""Host-side mirror of the reference's solver interface over the C ABI.

Names follow the reference: ``ProxADMMSolver(problem, params).solve()`` is
``ProxADMMSolver::Solve`` (src/epsilon/algorithms/prox_admm.cc:131-169) and what
``_solve.solve`` returns to Python (python/epopt/solvemodule.cc:110-186);
``eval_prox`` is ``_solve.eval_prox`` (solvemodule.cc:189-242); ``LinearMap.apply`` is
``LinearMapImpl::Apply`` (src/epsilon/linear/linear_map.h:52).  All arithmetic happens in
libepsgpu.so on the GPU; this module only marshals buffers.
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, Dict, List, Optional, Tuple

import numpy as np
import scipy.sparse as sp

from . import capi
from .problem import (MAP_DENSE, MAP_DIAGONAL, MAP_KRONECKER, MAP_SCALAR, MAP_SPARSE, PROX_TYPES,
                      Affine, LinearMapSpec, ProxAffineProblem, ProxFunction, Term, Variable)

STATE_NAMES = {0: "NOT_STARTED", 1: "INITIALIZING", 2: "RUNNING", 3: "OPTIMAL",
               4: "MAX_ITERATIONS_REACHED", 5: "ERROR"}
KKT_MODES = {"auto": 0, "diagonal": 1, "woodbury": 2, "gram": 3, "cg": 4}
# SolverParams.Solver (proto/epsilon/solver_params.proto:63-67)
SOLVERS = {"PROX_ADMM": 0, "PROX_ADMM_TWO_BLOCK": 1, "prox_admm": 0, "prox_admm_two_block": 1}

# SolverParams defaults (proto/epsilon/solver_params.proto:4-71)
PARAM_FIELDS = {"max_iterations", "rho", "abs_tol", "rel_tol", "epoch_iterations", "warm_start",
                "kkt_mode", "cg_tol", "cg_max_iterations", "use_graph", "profile", "gram_max_n", "solver"}


class Context:
    """One per process and GPU (eps_ctx)."""

    def __init__(self, device: int = 0, stream: int = 0):
        self.lib = capi.load()
        self.handle = C.c_void_p()
        capi.check(self.lib.eps_ctx_create(device, C.c_void_p(stream or None), C.byref(self.handle)))
        self.device = device
        self.n_ranks, self.rank = 1, 0

    def init_comm(self, n_ranks: int, rank: int, unique_id: bytes) -> None:
        buf = C.create_string_buffer(unique_id, 128)
        capi.check(self.lib.eps_ctx_init_comm(self.handle, n_ranks, rank, buf))
        self.n_ranks, self.rank = n_ranks, rank

    @staticmethod
    def nccl_unique_id() -> bytes:
        lib = capi.load()
        buf = C.create_string_buffer(128)
        capi.check(lib.eps_nccl_unique_id(buf))
        return buf.raw

    @property
    def kernel_launches(self) -> int:
        return int(self.lib.eps_ctx_kernel_launches(self.handle))

    @property
    def bytes_allocated(self) -> int:
        return int(self.lib.eps_ctx_bytes_allocated(self.handle))

    def synchronize(self) -> None:
        capi.check(self.lib.eps_ctx_synchronize(self.handle))

    def close(self) -> None:
        if self.handle:
            self.lib.eps_ctx_destroy(self.handle)
            self.handle = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


class LinearMap:
    """Device-resident LinearMapImpl (dense / sparse / diagonal / scalar / Kronecker)."""

    def __init__(self, ctx: Context, spec: Optional[LinearMapSpec] = None, _handle=None):
        self.ctx = ctx
        self.lib = ctx.lib
        self.handle = C.c_void_p()
        self._keep = []
        if _handle is not None:
            self.handle = _handle
            return
        h = C.byref(self.handle)
        if spec.kind == MAP_DENSE and spec.device_ptr is not None:
            self._keep = [spec.device_owner]
            capi.check(self.lib.eps_map_dense_dev(ctx.handle, spec.m, spec.n, C.c_void_p(spec.device_ptr), h))
        elif spec.kind == MAP_DENSE:
            a = np.asfortranarray(spec.dense, dtype=np.float64)
            capi.check(self.lib.eps_map_dense(ctx.handle, a.shape[0], a.shape[1], capi.dptr(a), h))
        elif spec.kind == MAP_SPARSE:
            # a CSC matrix is taken as it is: re-wrapping it would drop scipy's cached "indices are
            # sorted" flag and sort_indices() would rescan all nnz entries on the host (60 ms at nnz = 1e8)
            a = spec.sparse if sp.isspmatrix_csc(spec.sparse) else sp.csc_matrix(spec.sparse)
            a.sort_indices()
            indptr = np.ascontiguousarray(a.indptr, dtype=np.int32)
            indices = np.ascontiguousarray(a.indices, dtype=np.int32)
            data = np.ascontiguousarray(a.data, dtype=np.float64)
            capi.check(self.lib.eps_map_csc(ctx.handle, a.shape[0], a.shape[1], a.nnz,
                                            capi.iptr(indptr), capi.iptr(indices), capi.dptr(data), h))
        elif spec.kind == MAP_DIAGONAL:
            d = np.ascontiguousarray(spec.diag, dtype=np.float64)
            capi.check(self.lib.eps_map_diagonal(ctx.handle, d.size, capi.dptr(d), h))
        elif spec.kind == MAP_SCALAR:
            capi.check(self.lib.eps_map_scalar(ctx.handle, spec.n, float(spec.scalar), h))
        elif spec.kind == MAP_KRONECKER:
            a, b = LinearMap(ctx, spec.kron[0]), LinearMap(ctx, spec.kron[1])
            self._keep = [a, b]
            capi.check(self.lib.eps_map_kronecker(ctx.handle, a.handle, b.handle, h))
        else:
            raise ValueError("unknown map kind %r" % spec.kind)

    @property
    def shape(self) -> Tuple[int, int]:
        m, n = C.c_int(), C.c_int()
        capi.check(self.lib.eps_map_dims(self.handle, C.byref(m), C.byref(n)))
        return m.value, n.value

    def transpose(self) -> "LinearMap":
        h = C.c_void_p()
        capi.check(self.lib.eps_map_transpose(self.handle, C.byref(h)))
        t = LinearMap(self.ctx, _handle=h)
        t._keep = [self]
        return t

    def inverse(self) -> "LinearMap":
        """LinearMapImpl::Inverse() (linear_map.h:49)."""
        h = C.c_void_p()
        capi.check(self.lib.eps_map_inverse(self.handle, C.byref(h)))
        t = LinearMap(self.ctx, _handle=h)
        t._keep = [self]
        return t

    def apply(self, x: np.ndarray) -> np.ndarray:
        m, n = self.shape
        x = np.ascontiguousarray(x, dtype=np.float64).ravel()
        assert x.size == n, (x.size, n)
        y = np.empty(m, dtype=np.float64)
        capi.check(self.lib.eps_map_apply(self.handle, capi.dptr(x), capi.dptr(y)))
        return y

    def time(self, reps: int = 20) -> Tuple[float, float]:
        """(milliseconds per device apply by CUDA events, algorithmic bytes per apply)."""
        ms, nbytes = C.c_float(), C.c_double()
        capi.check(self.lib.eps_map_time(self.handle, reps, C.byref(ms), C.byref(nbytes)))
        return ms.value, nbytes.value

    def close(self) -> None:
        if self.handle:
            self.lib.eps_map_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class _ProblemBuilder:
    """ProxAffineProblem -> eps_problem (the AffineOperator terms of affine.cc:22-129)."""

    def __init__(self, ctx: Context, problem: ProxAffineProblem,
                 sharded_args: Optional[Dict[Tuple[int, int], int]] = None,
                 sharded_constraints: Optional[Dict[int, int]] = None):
        self.ctx, self.lib = ctx, ctx.lib
        self.handle = C.c_void_p()
        self.maps: List[LinearMap] = []
        self._map_cache: Dict[int, LinearMap] = {}
        capi.check(self.lib.eps_problem_create(ctx.handle, C.byref(self.handle)))
        self.var_index: Dict[str, int] = {}
        self.var_size: Dict[str, int] = {}
        for var_id, size in problem.variables().items():
            idx = C.c_int()
            capi.check(self.lib.eps_problem_add_variable(self.handle, var_id.encode(), size, C.byref(idx)))
            self.var_index[var_id] = idx.value
            self.var_size[var_id] = size
        for c in problem.constraints:
            ci = C.c_int()
            capi.check(self.lib.eps_problem_add_constraint(self.handle, c.rows, C.byref(ci)))
            if sharded_constraints and ci.value in sharded_constraints:
                capi.check(self.lib.eps_problem_constraint_sharded(self.handle, ci.value,
                                                                   int(sharded_constraints[ci.value])))
            for t in c.terms:
                self._add_term(-1, ci.value, t)
        for f in problem.prox:
            fi = C.c_int()
            capi.check(self.lib.eps_problem_add_prox(self.handle, PROX_TYPES[f.type], float(f.alpha),
                                                     int(f.epigraph), C.byref(fi)))
            if f.type == "SUM_DEADZONE" or f.type == "SUM_QUANTILE":
                qa = qb = None
                nq = 0
                if f.quantile_alpha is not None:
                    qa = np.ascontiguousarray(f.quantile_alpha, dtype=np.float64).ravel()
                    qb = np.ascontiguousarray(f.quantile_beta, dtype=np.float64).ravel()
                    nq = qa.size
                capi.check(self.lib.eps_problem_prox_zone(
                    self.handle, fi.value, float(f.zone_m),
                    capi.dptr(qa) if qa is not None else None,
                    capi.dptr(qb) if qb is not None else None, nq))
            if f.type == "SECOND_ORDER_CONE":
                capi.check(self.lib.eps_problem_prox_soc_dims(self.handle, fi.value, int(f.soc_dims[0]),
                                                              int(f.soc_dims[1])))
            if f.mat_dims is not None:
                capi.check(self.lib.eps_problem_prox_mat_dims(self.handle, fi.value, int(f.mat_dims[0]),
                                                              int(f.mat_dims[1])))
            if f.axis is not None:
                capi.check(self.lib.eps_problem_prox_axis(self.handle, fi.value, int(f.axis), int(f.axis_dims[0]),
                                                          int(f.axis_dims[1])))
            if f.type == "SUM_LARGEST":
                capi.check(self.lib.eps_problem_prox_sum_largest(self.handle, fi.value, int(f.sum_largest_k)))
            for ai_, a in enumerate(f.args):
                ai = C.c_int()
                capi.check(self.lib.eps_problem_prox_add_arg(self.handle, fi.value, a.rows, C.byref(ai)))
                if sharded_args and (fi.value, ai_) in sharded_args:
                    capi.check(self.lib.eps_problem_prox_arg_sharded(
                        self.handle, fi.value, ai.value, int(sharded_args[(fi.value, ai_)])))
                for t in a.terms:
                    self._add_term(fi.value, ai.value, t)

    def _map(self, spec: Optional[LinearMapSpec]):
        if spec is None:
            return None
        key = id(spec)
        if key not in self._map_cache:
            m = LinearMap(self.ctx, spec)
            self._map_cache[key] = m
            self.maps.append(m)
        return self._map_cache[key].handle

    def _add_term(self, prox_index: int, index: int, t: Term) -> None:
        mh = self._map(t.map)
        if isinstance(t.operand, Variable):
            capi.check(self.lib.eps_problem_add_term(self.handle, prox_index, index, mh,
                                                     self.var_index[t.operand.var_id], None))
        else:
            c = np.ascontiguousarray(t.operand, dtype=np.float64).ravel()
            capi.check(self.lib.eps_problem_add_term(self.handle, prox_index, index, mh, -1, capi.dptr(c)))

    def close(self) -> None:
        if self.handle:
            self.lib.eps_problem_destroy(self.handle)
            self.handle = C.c_void_p()
        for m in self.maps:
            m.close()
        self.maps = []


def make_params(lib, **kw) -> capi.Params:
    p = capi.Params()
    lib.eps_params_default(C.byref(p))
    for k, v in kw.items():
        if k not in PARAM_FIELDS:
            raise TypeError("unknown solver parameter %r" % k)
        if k == "kkt_mode" and isinstance(v, str):
            v = KKT_MODES[v]
        if k == "solver" and isinstance(v, str):
            v = SOLVERS[v]
        setattr(p, k, int(v) if isinstance(v, bool) else v)
    return p


class ProxADMMSolver:
    """ProxADMMSolver of the reference, iteration loop resident on one B200."""

    def __init__(self, problem: ProxAffineProblem, params: Optional[dict] = None,
                 ctx: Optional[Context] = None,
                 sharded_args: Optional[Dict[Tuple[int, int], int]] = None,
                 sharded_constraints: Optional[Dict[int, int]] = None):
        self._own_ctx = ctx is None
        self.ctx = ctx or Context()
        self.lib = self.ctx.lib
        self.problem = problem
        self.params = make_params(self.lib, **(params or {}))
        self._builder = _ProblemBuilder(self.ctx, problem, sharded_args, sharded_constraints)
        self.handle = C.c_void_p()
        capi.check(self.lib.eps_solver_create(self._builder.handle, C.byref(self.params),
                                              C.byref(self.handle)))
        self.status = capi.Status()
        self.trace: List[dict] = []

    def describe(self) -> str:
        return self.lib.eps_solver_describe(self.handle).decode()

    def solve(self, callback: Optional[Callable[[dict], bool]] = None) -> Dict[str, np.ndarray]:
        """Run to OPTIMAL / max_iterations; returns {var_id: value} (GetSolution)."""
        self.trace = []

        def _cb(st, _user):
            d = st.contents.as_dict()
            self.trace.append(d)
            return 1 if (callback and callback(d)) else 0

        cb = capi.TRACE_FN(_cb)
        capi.check(self.lib.eps_solver_solve(self.handle, C.byref(self.status), cb, None))
        return self.solution()

    def set_constant(self, prox_index: int, index: int, values: np.ndarray) -> None:
        """Solver::SetParameterValue for a constant vector (solver.cc:109-116): the constant term of
        constraint `index` (prox_index < 0) or of argument `index` of prox function `prox_index`."""
        v = np.ascontiguousarray(values, dtype=np.float64).ravel()
        capi.check(self.lib.eps_solver_set_constant(self.handle, prox_index, index, capi.dptr(v)))

    def iterate(self, iterations: int) -> float:
        """Exactly `iterations` more ADMM iterations, no stop test; CUDA-event milliseconds."""
        ms = C.c_float()
        capi.check(self.lib.eps_solver_iterate(self.handle, iterations, C.byref(ms)))
        return ms.value

    def residuals(self) -> dict:
        st = capi.Status()
        capi.check(self.lib.eps_solver_residuals(self.handle, C.byref(st)))
        return st.as_dict()

    def solution(self) -> Dict[str, np.ndarray]:
        out = {}
        for var_id, size in self._builder.var_size.items():
            buf = np.empty(size, dtype=np.float64)
            capi.check(self.lib.eps_solver_get_variable(self.handle, var_id.encode(), capi.dptr(buf)))
            out[var_id] = buf
        return out

    def profile(self) -> Dict[str, dict]:
        n = 16
        names = (C.c_char_p * n)()
        launches = (C.c_int64 * n)()
        ms = (C.c_double * n)()
        nbytes = (C.c_double * n)()
        k = self.lib.eps_solver_profile(self.handle, n, names, launches, ms, nbytes)
        return {names[i].decode(): {"launches": int(launches[i]), "ms": float(ms[i]),
                                    "bytes": float(nbytes[i])} for i in range(k)}

    def close(self) -> None:
        if self.handle:
            self.lib.eps_solver_destroy(self.handle)
            self.handle = C.c_void_p()
        self._builder.close()
        if self._own_ctx:
            self.ctx.close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def solve(problem: ProxAffineProblem, ctx: Optional[Context] = None, **params):
    """`_solve.solve` analogue: returns (status dict, {var_id: value})."""
    with ProxADMMSolver(problem, params, ctx=ctx) as s:
        x = s.solve()
        st = s.status.as_dict()
        st["trace"] = list(s.trace)
        return st, x


def eval_prox(f: ProxFunction, lam: float, v: Dict[str, np.ndarray],
              ctx: Optional[Context] = None) -> Dict[str, np.ndarray]:
    """`_solve.eval_prox` analogue (solvemodule.cc:189-242): argmin lam*f(x) + 0.5||x - v||^2."""
    own = ctx is None
    ctx = ctx or Context()
    try:
        prob = ProxAffineProblem([f], [])
        b = _ProblemBuilder(ctx, prob)
        try:
            order = sorted(b.var_index, key=lambda k: b.var_index[k])
            vin = np.concatenate([np.asarray(v[k], dtype=np.float64).ravel() for k in order])
            xout = np.empty_like(vin)
            capi.check(ctx.lib.eps_prox_eval(b.handle, float(lam), capi.dptr(vin), capi.dptr(xout)))
            out, off = {}, 0
            for k in order:
                out[k] = xout[off:off + b.var_size[k]].copy()
                off += b.var_size[k]
            return out
        finally:
            b.close()
    finally:
        if own:
            ctx.close()
