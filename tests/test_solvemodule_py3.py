"""`_solve` for CPython 3 (adapter/solvemodule_py3.cc -> oracle/_ref/_solve.so): the binding python/epopt calls
(`_solve.solve(problem_bytes, parameters, params_bytes, data)`, `_solve.eval_prox(...)`, `_solve.error`;
python/epopt/solvemodule.cc:110-242,282-289), bytes in / bytes out, with ProxADMMGpuSolver behind it.

CPU part: with EPSILON_GPU=0 the module runs the STOCK reference classes, which pins the byte-level contract
(serialized Problem / SolverParams / SolverStatus, DataMap strings, raw f64 results) against the reference's
fixtures without a GPU.  GPU part: the same calls land on the device."""
import json
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")
sys.path.insert(0, GOLDEN)

import make_golden  # noqa: E402
from oracle import ref_proto  # noqa: E402

INDEX = json.load(open(os.path.join(GOLDEN, "index.json")))
TRACE_KEYS = make_golden.TRACE_KEYS


@pytest.fixture(scope="module")
def solve_mod():
    if not ref_proto.available():
        pytest.skip("oracle/_ref/_solve.so not built (make -C oracle py3; needs /root/reference)")
    return ref_proto.load_module()


def _solve(mod, name, extra_params=None, parameters=(), extra_data=None):
    prob, info, params = make_golden.make_problem(name)
    if make_golden.problem_digest(prob) != INDEX["solve"][name]["digest"]:
        pytest.skip("generator drift on this host")
    pb, parb, data = ref_proto.solve_inputs(prob, dict(params, **(extra_params or {})))
    data.update(extra_data or {})
    st_b, xs = mod.solve(pb, list(parameters), parb, data)
    st = ref_proto.messages()["SolverStatus"].FromString(st_b)
    x = {k: np.frombuffer(v, dtype=np.float64) for k, v in xs.items()}
    return prob, info, st, x


def _check(name, st, x, info, rtol):
    gold = np.load(os.path.join(GOLDEN, name + ".npz"))
    g = dict(zip(TRACE_KEYS, gold["status"]))
    assert st.state == int(g["state"]) and st.num_iterations == int(g["num_iterations"])
    for k in TRACE_KEYS[2:]:
        assert getattr(st.residuals, k) == pytest.approx(g[k], rel=rtol, abs=1e-12), k
    v = INDEX["solve"][name]["solution_var"]
    assert np.abs(x[v] - gold["x"]).max() <= 1e-7 * max(1.0, np.abs(gold["x"]).max())
    assert info["objective"](x) == pytest.approx(float(gold["objective"]), rel=1e-6)


@pytest.mark.parametrize("name", ["lasso_dense_100x500", "logreg_l1_2000x200", "tv_1d_5000"])
def test_module_contract_with_the_stock_solver(solve_mod, name, monkeypatch):
    monkeypatch.setenv("EPSILON_GPU", "0")
    prob, info, st, x = _solve(solve_mod, name)
    _check(name, st, x, info, rtol=1e-12)
    assert sorted(x) == sorted(prob.variables())       # every variable of the compiled problem comes back


def test_failed_check_is_a_python_exception(solve_mod, monkeypatch):
    """CHECK_EQ(1, rho) (prox_admm.cc:50) -> glog failure function -> longjmp -> _solve.error (solvemodule.cc:245-263)."""
    monkeypatch.setenv("EPSILON_GPU", "0")
    with pytest.raises(solve_mod.error):
        _solve(solve_mod, "lasso_dense_100x500", extra_params={"rho": 2.0})


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["lasso_dense_100x500", "lasso_sparse_4000x400", "logreg_l1_2000x200",
                                  "group_lasso_400x10x12", "tv_1d_5000", "weighted_lasso_200x300"])
def test_module_on_the_device(solve_mod, name):
    prob, info, st, x = _solve(solve_mod, name)
    _check(name, st, x, info, rtol=1e-7)


@pytest.mark.gpu
def test_module_device_errors_are_python_exceptions(solve_mod):
    with pytest.raises(solve_mod.error):
        _solve(solve_mod, "lasso_dense_100x500", extra_params={"rho": 2.0})


@pytest.mark.gpu
def test_module_warm_start_with_a_parameter(solve_mod, ref):
    """warm_start + parameters: the cached device solver gets Solver::SetParameterValue and continues from its
    iterate (solvemodule.cc:142-156); must equal the stock solver doing the same (harness mode `resolve`)."""
    from epsilon_b200 import problems
    rng = np.random.default_rng(5)
    prob, info = problems.lasso(100, 500)
    prob.prox[0].args[0].terms[1].parameter_id = "b"
    b2 = info["b"] + 0.2 * rng.standard_normal(100)
    params = {"max_iterations": 300, "warm_start": True}
    cpu = ref.run(prob, params={"max_iterations": 300}, mode="resolve", reparam={"b": b2})
    pb, parb, data = ref_proto.solve_inputs(prob, params)
    M = ref_proto.messages()
    st1_b, x1 = solve_mod.solve(pb, [], parb, data)
    c = M["Constant"](constant_type=1, m=100, n=1, data_location="param_b_new", parameter_id="b")
    st2_b, x2 = solve_mod.solve(pb, [("b", c.SerializeToString())], parb, {"param_b_new": b2.tobytes()})
    for st_b, want in ((st1_b, cpu["status_first"]), (st2_b, cpu["status"])):
        st = M["SolverStatus"].FromString(st_b)
        assert st.num_iterations == want["num_iterations"] and st.state == want["state"]
        for k in TRACE_KEYS[2:]:
            assert getattr(st.residuals, k) == pytest.approx(want[k], rel=1e-7, abs=1e-12)
    assert np.abs(np.frombuffer(x2["x"]) - cpu["x"]["x"]).max() <= 1e-7
    assert np.abs(np.frombuffer(x1["x"]) - cpu["x_first"]["x"]).max() <= 1e-7


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["norm_1", "sum_logistic", "norm_2_axis1", "norm_1_diag", "sum_square_tall"])
def test_module_eval_prox_on_the_device(solve_mod, name):
    f, lam, v = make_golden.prox_cases()[name]
    gold = np.load(os.path.join(GOLDEN, "prox_" + name + ".npz"))
    eb, data = ref_proto.prox_inputs(f, lam)
    out = solve_mod.eval_prox(eb, float(lam), data, {k: np.ascontiguousarray(val).tobytes() for k, val in v.items()})
    for k in v:
        got = np.frombuffer(out[k], dtype=np.float64)
        assert np.abs(got - gold["x_" + k]).max() <= 1e-10 * max(1.0, np.abs(gold["x_" + k]).max())
