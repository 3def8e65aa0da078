#!/usr/bin/env python
"""Roofline evidence for the fused elementwise prox kernels at an HBM-bound size (SURVEY.md section 7.3:
fused norm1 prox, 1e7 elements, >= 70 % of the measured HBM peak on 40*L algorithmic bytes).

Two denoising-shaped problems whose elementwise stage has L elements, run through eps_solver_iterate with the
per-class CUDA-event profile:   NORM_1:  SUM_SQUARE(xs - b) + NORM_1[lam](x),  xs - x = 0
                                LOGISTIC: SUM_LOGISTIC(t) + SUM_SQUARE(xs - b), t - xs = 0
  python tools/bench_prox.py [--n 10000000]        (one JSON line)
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from epsilon_b200.problem import Affine, LinearMapSpec, ProxAffineProblem, ProxFunction, Term, Variable  # noqa: E402
from epsilon_b200.solver import Context, ProxADMMSolver  # noqa: E402


def problem(kind, n, rng):
    b = rng.standard_normal(n) * 2
    xs, z = Variable("xs", n), Variable("z", n)
    f0 = ProxFunction("SUM_SQUARE", alpha=0.5, args=[Affine(n, [Term(xs), Term(b, LinearMapSpec.Scalar(-1.0, n), name="b")])])
    f1 = ProxFunction("NORM_1", alpha=0.7, args=[Affine(n, [Term(z)])]) if kind == "NORM_1" else \
        ProxFunction("SUM_LOGISTIC", args=[Affine(n, [Term(z)])])
    c0 = Affine(n, [Term(xs), Term(z, LinearMapSpec.Scalar(-1.0, n))])
    return ProxAffineProblem([f0, f1], [c0], name="%s_denoise_%d" % (kind.lower(), n))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=10_000_000)
    ap.add_argument("--steps", type=int, default=20)
    a = ap.parse_args()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    ctx = Context(0)
    out = {"n": a.n, "hbm_peak_GBps": peak, "kernels": {}}
    rng = np.random.default_rng(0)
    for kind in ("NORM_1", "SUM_LOGISTIC"):
        prob = problem(kind, a.n, rng)
        with ProxADMMSolver(prob, {"profile": 1, "use_graph": 0}, ctx=ctx) as s:
            s.iterate(3)
            base = s.profile()
            s.iterate(a.steps)
            prof = s.profile()
        v, b = prof["prox"], base["prox"]
        l, t, by = v["launches"] - b["launches"], v["ms"] - b["ms"], v["bytes"] - b["bytes"]
        out["kernels"]["ElemProxVec2Kernel<%s>" % kind] = {
            "launches": l, "us_per_launch": 1e3 * t / l, "algorithmic_bytes_per_launch": by / l,
            "GBps": by / (t / 1e3) / 1e9, "frac_of_measured_hbm_peak": by / (t / 1e3) / 1e9 / peak}
    ctx.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
