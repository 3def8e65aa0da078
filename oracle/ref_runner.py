"""Run the compiled reference (oracle/_ref/epsilon_ref) on a ProxAffineProblem.

TEST INFRASTRUCTURE -- only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this.  The product
(epsilon_b200/) never does.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import tempfile
from typing import Dict, Optional

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
BINARY = os.path.join(HERE, "_ref", "epsilon_ref")
# the same harness with adapter/prox_admm_gpu.cc (the reference-side binding of libepsgpu.so) linked in
BINARY_GPU = os.path.join(HERE, "_ref", "epsilon_ref_gpu")

STATE_NAMES = {0: "NOT_STARTED", 1: "INITIALIZING", 2: "RUNNING", 3: "OPTIMAL",
               4: "MAX_ITERATIONS_REACHED", 5: "ERROR"}


def available() -> bool:
    return os.path.isfile(BINARY) and os.access(BINARY, os.X_OK)


def gpu_adapter_available() -> bool:
    return os.path.isfile(BINARY_GPU) and os.access(BINARY_GPU, os.X_OK)


def _parse_status_line(line: str) -> dict:
    f = line.split()
    return {"state": int(f[0]), "num_iterations": int(f[1]), "r_norm": float(f[2]),
            "s_norm": float(f[3]), "epsilon_primal": float(f[4]), "epsilon_dual": float(f[5])}


def _read_vectors(path: str) -> Dict[str, np.ndarray]:
    out = {}
    if os.path.isdir(path):
        for name in os.listdir(path):
            out[name] = np.fromfile(os.path.join(path, name), dtype=np.float64)
    return out


def run(problem, params: Optional[dict] = None, mode: str = "trace", dump_iter: int = -1,
        workdir: Optional[str] = None, timeout: float = 3600.0, reparam: Optional[dict] = None,
        env_extra: Optional[dict] = None) -> dict:
    """mode 'solve' = stock ProxADMMSolver::Solve(); 'trace' = stepped, per-epoch residuals;
    'resolve' = Solve, SetParameterValue(reparam), Solve again (warm start); 'gpu-solve' / 'gpu-resolve' =
    the same flows through ProxADMMGpuSolver (adapter/prox_admm_gpu.cc over libepsgpu.so)."""
    from oracle.ref_format import write_reference_problem
    gpu = mode.startswith("gpu-")
    binary = BINARY_GPU if gpu else BINARY
    assert os.path.isfile(binary), "%s missing: run `make -C oracle%s`" % (binary, " gpu" if gpu else "")
    tmp = workdir or tempfile.mkdtemp(prefix="epsref_")
    try:
        write_reference_problem(problem, tmp, params, reparam)
        cmd = [binary, mode, tmp] + ([str(dump_iter)] if mode == "trace" else [])
        env = dict(os.environ, OMP_NUM_THREADS="1")
        env.update(env_extra or {})
        subprocess.run(cmd, check=True, timeout=timeout, env=env)
        out = os.path.join(tmp, "out_" + mode)
        lines = open(os.path.join(out, "status.txt")).read().splitlines()
        res = {"status": _parse_status_line(lines[0])}
        for ln in lines[1:]:
            k, v = ln.split()
            res[k] = float(v)
        tr = os.path.join(out, "trace.txt")
        if os.path.exists(tr):
            res["trace"] = [_parse_status_line(l) for l in open(tr).read().splitlines() if l.strip()]
        res["x"] = _read_vectors(os.path.join(out, "x"))
        for extra in ("trace_first", ):
            trf = os.path.join(out, extra + ".txt")
            if os.path.exists(trf):
                res[extra] = [_parse_status_line(l) for l in open(trf).read().splitlines() if l.strip()]
        sf = os.path.join(out, "status_first.txt")
        if os.path.exists(sf):
            res["status_first"] = _parse_status_line(open(sf).read().splitlines()[0])
            res["x_first"] = _read_vectors(os.path.join(out, "x_first"))
        df = os.path.join(out, "describe.txt")
        if os.path.exists(df):
            res["describe"] = open(df).read()
        if dump_iter >= 0:
            res["apply"] = {}
            for name in sorted(os.listdir(out)):
                if name.startswith("apply_"):
                    res["apply"][name] = _read_vectors(os.path.join(out, name))
        return res
    finally:
        if workdir is None:
            shutil.rmtree(tmp, ignore_errors=True)


def run_prox(f, lam: float, v: Dict[str, np.ndarray], workdir: Optional[str] = None) -> Dict[str, np.ndarray]:
    """Single operator through the EvalProx statement sequence (solvemodule.cc:189-242)."""
    from oracle.ref_format import write_reference_prox
    assert available(), "oracle/_ref/epsilon_ref missing: run `make -C oracle`"
    tmp = workdir or tempfile.mkdtemp(prefix="epsref_")
    try:
        write_reference_prox(f, lam, v, tmp)
        subprocess.run([BINARY, "prox", tmp], check=True, timeout=3600)
        return _read_vectors(os.path.join(tmp, "out_prox", "x"))
    finally:
        if workdir is None:
            shutil.rmtree(tmp, ignore_errors=True)
