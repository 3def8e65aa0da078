#!/usr/bin/env python
"""Generate tests/golden/lasso_sparse_1Mx100k.npz: the pinned CPU oracle on the exact bench instance.

BASELINE config 2 (sparse lasso 1M x 100k, density 1e-3) is the instance bench.py times and the
one whose kernels (SellStreamKernel + the fused CG tail) produce every headline number.  The
compiled reference cannot run it (it densifies and inverts a 100k x 100k matrix,
src/epsilon/linear/sparse_matrix_impl.cc:60-78), so the fixture comes from the numpy restatement
oracle/prox_admm_np.py -- itself pinned to the reference's outputs by tests/test_oracle_golden.py --
with its KKT prox solved matrix-free ("kkt": "cg", conjugate gradients to 1e-13).  Everything else
(prox_admm.cc:131-217: the iteration, ComputeResiduals, the stop test) is the same code that the
small fixtures hold to the reference.

    python tests/golden/make_fullsize.py            # ~10 min, ~12 GB of host memory

Recorded: per-epoch {state, num_iterations, r, s, eps_pri, eps_dual}, the final status, the solution
x (sparse, so the compressed file is small), the objective, SHA-256 of x and of the generated inputs.
"""
from __future__ import annotations

import hashlib
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from epsilon_b200 import problems  # noqa: E402
from oracle import prox_admm_np as onp  # noqa: E402

M, N, DENSITY, PANELS, SEED = 1_000_000, 100_000, 1e-3, 8, 0
TRACE_KEYS = ("state", "num_iterations", "r_norm", "s_norm", "epsilon_primal", "epsilon_dual")
NAME = "lasso_sparse_1Mx100k"


def input_digest(info) -> str:
    a = info["A"]
    h = hashlib.sha256()
    h.update(a.indptr.astype(np.int32).tobytes())
    h.update(a.indices.astype(np.int32).tobytes())
    h.update(a.data.astype(np.float64).tobytes())
    h.update(np.asarray(info["b"], dtype=np.float64).tobytes())
    h.update(np.float64(info["lam"]).tobytes())
    return h.hexdigest()


def main(m=M, n=N, density=DENSITY, name=NAME):
    t0 = time.time()
    prob, info = problems.lasso_sparse_panels(m, n, density, panels=PANELS, seed=SEED)
    print("generated in %.1f s, nnz = %d" % (time.time() - t0, info["A"].nnz), flush=True)
    digest = input_digest(info)
    t0 = time.time()
    o = onp.ProxADMM(prob, {"kkt": "cg"})
    print("oracle init %.1f s" % (time.time() - t0), flush=True)
    t0 = time.time()
    r = o.solve()
    cg = sum(getattr(op, "cg_iterations", 0) for op in o.ops)
    print("oracle solve %.1f s: %s, %d CG steps" % (time.time() - t0, r["status"], cg), flush=True)
    x = r["x"]["x"]
    trace = np.array([[t[k] for k in TRACE_KEYS] for t in r["trace"]], dtype=np.float64)
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"), trace=trace,
        status=np.array([r["status"][k] for k in TRACE_KEYS], dtype=np.float64),
        mnN=np.array([r["m"], r["n"], r["N"]], dtype=np.int64), x=x,
        objective=np.float64(info["objective"](r["x"])), lam=np.float64(info["lam"]),
        cg_iterations=np.int64(cg))
    meta = {"digest": digest, "x_sha256": hashlib.sha256(x.tobytes()).hexdigest(), "m": m, "n": n,
            "density": density, "panels": PANELS, "seed": SEED, "nnz": int(info["A"].nnz),
            "epochs": int(trace.shape[0]), "num_iterations": int(r["status"]["num_iterations"]),
            "state": int(r["status"]["state"]), "objective": float(info["objective"](r["x"])),
            "oracle": "oracle/prox_admm_np.py ProxADMM(kkt='cg', tol 1e-13)", "cg_iterations": int(cg)}
    with open(os.path.join(HERE, name + ".json"), "w") as fh:
        json.dump(meta, fh, indent=1, sort_keys=True)
    print(json.dumps(meta))


if __name__ == "__main__":
    main()
