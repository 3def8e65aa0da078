"""Where the end-to-end solve time goes (bench.py's e2e leg, one rank): Python-side wall clock around
problem construction (uploads, transposition, SELL build) and solve (Init + iterations), with the
library's own EPS_TIMING=1 phase lines on stderr.  Usage: EPS_TIMING=1 python tools/e2e_phases.py"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from epsilon_b200 import problems  # noqa: E402
from epsilon_b200.solver import Context, ProxADMMSolver  # noqa: E402


def main():
    m = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    n, density = 100_000, 1e-3
    prob, info = problems.lasso_sparse_panels(m, n, density, panels=8, my_panels=range(8), seed=0,
                                              allreduce=lambda a: a)
    a = info["A"]
    a.sort_indices()
    pinned = [torch.from_numpy(v).pin_memory() for v in (a.data, a.indices, a.indptr)]
    a.data, a.indices, a.indptr = (t.numpy() for t in pinned)
    ctx = Context(0)
    out = []
    for rep in range(int(sys.argv[2]) if len(sys.argv) > 2 else 4):
        ctx.synchronize()
        t0 = time.time()
        s = ProxADMMSolver(prob, {"cg_tol": 1e-12}, ctx=ctx)
        ctx.synchronize()
        t1 = time.time()
        x = s.solve()
        ctx.synchronize()
        t2 = time.time()
        st = s.status.as_dict()
        s.close()
        ctx.synchronize()
        t3 = time.time()
        out.append({"construct_s": t1 - t0, "solve_s": t2 - t1, "close_s": t3 - t2, "init_s": st["init_ms"] / 1e3,
                    "loop_s": st["loop_ms"] / 1e3, "iterations": st["iterations_run"]})
        print(json.dumps(out[-1]), flush=True)


if __name__ == "__main__":
    main()
