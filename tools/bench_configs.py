#!/usr/bin/env python
"""Per-config measurements on ONE GPU for the BASELINE.json configs other than the headline
one (bench.py covers config 2).  Prints one JSON line per config: iterations/s of the ADMM
loop, time-to-tolerance of a whole solve, and CUDA-event time / algorithmic bytes per kernel
class from a profiled pass, against the measured HBM peak.

  python tools/bench_configs.py tv1d --n 10000000
  python tools/bench_configs.py logreg --m 2000000 --n 200000 --density 1e-3
  python tools/bench_configs.py hinge --m 100000 --n 10000
  python tools/bench_configs.py group_lasso --m 50000 --group 50 --K 200
  python tools/bench_configs.py lasso_dense --m 1000 --n 5000
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config", choices=["tv1d", "logreg", "hinge", "group_lasso", "lasso_dense", "lasso_sparse", "mv_lasso"])
    ap.add_argument("--m", type=int, default=0)
    ap.add_argument("--n", type=int, default=0)
    ap.add_argument("--density", type=float, default=1.0)
    ap.add_argument("--group", type=int, default=50)
    ap.add_argument("--K", type=int, default=100)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--kkt-mode", default="auto")
    ap.add_argument("--max-iterations", type=int, default=2000)
    ap.add_argument("--no-solve", action="store_true")
    a = ap.parse_args()

    from epsilon_b200 import problems
    from epsilon_b200.solver import Context, ProxADMMSolver

    t0 = time.time()
    if a.config == "tv1d":
        prob, info = problems.tv_1d(a.n or 10_000_000, seed=0)
    elif a.config == "logreg":
        prob, info = problems.logreg_l1(a.m, a.n, a.density, seed=0)
    elif a.config == "hinge":
        prob, info = problems.hinge_l2(a.m, a.n, a.density, seed=0)
    elif a.config == "group_lasso":
        prob, info = problems.group_lasso(a.m, a.group, a.K, seed=0)
    elif a.config == "mv_lasso":
        prob, info = problems.mv_lasso(a.m, a.n, a.K, a.density, seed=0)
    elif a.config == "lasso_dense":
        prob, info = problems.lasso(a.m, a.n, 1.0, seed=0)
    else:
        prob, info = problems.lasso(a.m, a.n, a.density, seed=0)
    t_gen = time.time() - t0

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))

    ctx = Context(0)
    params = {"kkt_mode": a.kkt_mode, "max_iterations": a.max_iterations}
    out = {"config": a.config, "problem": prob.name, "data_generation_s": round(t_gen, 2)}

    if not a.no_solve:
        t0 = time.time()
        s = ProxADMMSolver(prob, params, ctx=ctx)
        t_init = time.time() - t0
        x = s.solve()
        ctx.synchronize()
        t_all = time.time() - t0
        st = s.status.as_dict()
        out["solve"] = {"state": st["state"], "num_iterations": st["num_iterations"],
                        "iterations_run": st["iterations_run"], "init_s": round(t_init, 4),
                        "loop_s": st["loop_ms"] / 1e3, "time_to_tol_s": round(t_all, 4),
                        "objective": info["objective"](x), "r_norm": st["r_norm"], "s_norm": st["s_norm"],
                        "cg_iterations": st["cg_iterations"]}
        out["solver"] = s.describe().strip().split("\n")
        s.close()

    s = ProxADMMSolver(prob, params, ctx=ctx)
    s.iterate(a.warmup)
    ms = s.iterate(a.steps)
    out["iterations_per_s"] = a.steps / (ms / 1e3)
    out["ms_per_iteration"] = ms / a.steps
    s.close()

    s = ProxADMMSolver(prob, dict(params, profile=1), ctx=ctx)
    s.iterate(a.warmup)
    base = s.profile()
    s.iterate(a.steps)
    prof = s.profile()
    classes = {}
    for k, v in prof.items():
        b = base.get(k, {"launches": 0, "ms": 0.0, "bytes": 0.0})
        l, t, by = v["launches"] - b["launches"], v["ms"] - b["ms"], v["bytes"] - b["bytes"]
        if l > 0:
            classes[k] = {"launches": l, "ms_per_iteration": t / a.steps, "us_per_launch": 1e3 * t / l,
                          "GBps": by / (t / 1e3) / 1e9 if t > 0 else None,
                          "frac_of_measured_hbm_peak": by / (t / 1e3) / 1e9 / peak if t > 0 else None}
    out["classes"] = classes
    out["hbm_peak_GBps"] = peak
    s.close()
    ctx.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
