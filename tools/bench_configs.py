#!/usr/bin/env python
"""Per-config measurements for the BASELINE.json configs other than the headline one (bench.py covers
config 2).  One JSON line per run: iterations/s of the ADMM loop, time-to-tolerance of a whole solve, and
CUDA-event time / algorithmic bytes per kernel class from a profiled pass, against the measured HBM peak.

One GPU:
  python tools/bench_configs.py tv1d --n 10000000
  python tools/bench_configs.py logreg --m 2000000 --n 200000 --density 1e-3
  python tools/bench_configs.py hinge --m 1000000 --n 10000 --device-gen        # 80 GB matrix made on the GPU
  python tools/bench_configs.py group_lasso --m 50000 --group 50 --K 200
  python tools/bench_configs.py lasso_dense --m 1000 --n 5000
N GPUs (sample rows of the constraint block sharded, eps_problem_constraint_sharded):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \\
      tools/bench_configs.py logreg --rows 2000000 --cols 200000 --density 1e-3
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


def hinge_l2_on_device(m, n, rank, world, seed, allreduce_dev, lam=1.0):
    """hinge_l2 (problems/hinge_l2.py: dense N(0,1) data, unit-norm columns, y = sign(A x0 + noise)) with the data
    block generated ON THE DEVICE by torch's counter-based (Philox) generator, in the row-sharded compiled form:
    this rank holds rows [rank*m/world, (rank+1)*m/world).  The 8*m*n-byte matrix never exists on the host and
    is handed to the library without a copy (eps_map_dense_dev)."""
    import torch
    from epsilon_b200.problem import Affine, LinearMapSpec, ProxAffineProblem, ProxFunction, Term, Variable
    lo, hi = rank * m // world, (rank + 1) * m // world
    ml = hi - lo
    dev = torch.device("cuda", torch.cuda.current_device())
    g = torch.Generator(device=dev)
    g.manual_seed(seed * 1000 + rank)
    a = torch.empty((n, ml), dtype=torch.float64, device=dev)       # row j = column j of the column-major ml x n block
    step = max(1, (1 << 28) // max(ml, 1))
    sq = torch.empty(n, dtype=torch.float64, device=dev)
    for j in range(0, n, step):
        blk = a[j:j + step]
        blk.normal_(generator=g)
        sq[j:j + step] = (blk * blk).sum(dim=1)
    sq = allreduce_dev(sq)
    inv = 1.0 / torch.sqrt(sq)
    x0 = np.zeros(n)
    rng = np.random.default_rng([seed, 10_000])
    k = max(1, int(round(0.01 * n)))
    idx = rng.choice(n, size=k, replace=False)
    x0[idx] = rng.standard_normal(k)
    x0d = torch.from_numpy(x0).to(dev)
    z = torch.zeros(ml, dtype=torch.float64, device=dev)
    for j in range(0, n, step):
        blk = a[j:j + step]
        blk *= inv[j:j + step, None]
        z += x0d[j:j + step] @ blk
    z += 0.05 * torch.randn(ml, dtype=torch.float64, device=dev, generator=g)
    y = torch.sign(z)
    y[y == 0] = 1.0
    for j in range(0, n, step):
        a[j:j + step] *= -y[None, :]
    torch.cuda.synchronize()
    nda = LinearMapSpec(0, ml, n, name="negDA", device_ptr=a.data_ptr(), device_owner=a)
    t, x = Variable("t", ml), Variable("x", n)
    f0 = ProxFunction("SUM_HINGE", args=[Affine(ml, [Term(t)])])
    f1 = ProxFunction("SUM_SQUARE", alpha=lam, args=[Affine(n, [Term(x)])])
    c0 = Affine(ml, [Term(np.ones(ml), name="ones"), Term(x, nda), Term(t, LinearMapSpec.Scalar(-1.0, ml))])

    def objective(sol):
        xv = torch.from_numpy(np.ascontiguousarray(sol["x"])).to(dev)
        zz = torch.zeros(ml, dtype=torch.float64, device=dev)
        for j in range(0, n, step):
            zz += xv[j:j + step] @ a[j:j + step]
        part = torch.clamp(1.0 + zz, min=0).sum().reshape(1)
        return float(allreduce_dev(part)[0]) + lam * float(sol["x"] @ sol["x"])

    return ProxAffineProblem([f0, f1], [c0], name="hinge_l2_%dx%d_device" % (m, n)), {
        "objective": objective, "solution_var": "x", "lam": lam, "sharded_constraints": {0: m}, "keep": a}


def group_lasso_on_device(m, gs, K, rank, world, seed, allreduce_dev):
    """group_lasso (problems/group_lasso.py with K equal groups of `gs`, as BASELINE config 4 asks) with the dense
    m x n data matrix generated on the device, rows sharded over the ranks (the 500k x 50k matrix is 200 GB: it
    does not fit one GPU).  Compiled form of epsilon_b200.problems.group_lasso; the SUM_SQUARE argument is the
    row-sharded one (eps_problem_prox_arg_sharded)."""
    import scipy.sparse as sp
    import torch
    from epsilon_b200.problem import Affine, LinearMapSpec, ProxAffineProblem, ProxFunction, Term, Variable
    n = gs * K
    lo, hi = rank * m // world, (rank + 1) * m // world
    ml = hi - lo
    dev = torch.device("cuda", torch.cuda.current_device())
    g = torch.Generator(device=dev)
    g.manual_seed(seed * 1000 + rank)
    a = torch.empty((n, ml), dtype=torch.float64, device=dev)
    step = max(1, (1 << 28) // max(ml, 1))
    sq = torch.empty(n, dtype=torch.float64, device=dev)
    for j in range(0, n, step):
        blk = a[j:j + step]
        blk.normal_(generator=g)
        sq[j:j + step] = (blk * blk).sum(dim=1)
    inv = 1.0 / torch.sqrt(allreduce_dev(sq))
    rng = np.random.default_rng([seed, 10_000])
    x0 = np.zeros(n)
    for k in range(K):
        if rng.random() < 0.2:
            x0[k * gs:(k + 1) * gs] = rng.standard_normal(gs)
    x0d = torch.from_numpy(x0).to(dev)
    b = torch.zeros(ml, dtype=torch.float64, device=dev)
    for j in range(0, n, step):
        blk = a[j:j + step]
        blk *= inv[j:j + step, None]
        b += x0d[j:j + step] @ blk
    b += np.sqrt(0.001) * torch.randn(ml, dtype=torch.float64, device=dev, generator=g)
    atb = torch.empty(n, dtype=torch.float64, device=dev)
    for j in range(0, n, step):
        atb[j:j + step] = a[j:j + step] @ b
    atb = allreduce_dev(atb).cpu().numpy()
    lam = 0.1 * max(float(np.linalg.norm(atb[k * gs:(k + 1) * gs])) for k in range(K))
    torch.cuda.synchronize()
    x = Variable("x", n)
    amap = LinearMapSpec(0, ml, n, name="A", device_ptr=a.data_ptr(), device_owner=a)
    bh = b.cpu().numpy()
    prox = [ProxFunction("SUM_SQUARE", alpha=0.5, args=[Affine(ml, [Term(x, amap), Term(bh, LinearMapSpec.Scalar(-1.0, ml), name="b")])])]
    cons = []
    for k in range(K):
        tk = Variable("t%04d" % k, gs)
        prox.append(ProxFunction("NORM_2", alpha=lam, args=[Affine(gs, [Term(tk)])]))
        sel = sp.csc_matrix((np.ones(gs), (np.arange(gs), np.arange(k * gs, (k + 1) * gs))), shape=(gs, n))
        cons.append(Affine(gs, [Term(x, LinearMapSpec.Sparse(sel, name="S%d" % k)), Term(tk, LinearMapSpec.Scalar(-1.0, gs))]))

    def objective(sol):
        xv = torch.from_numpy(np.ascontiguousarray(sol["x"])).to(dev)
        r = -b.clone()
        for j in range(0, n, step):
            r += xv[j:j + step] @ a[j:j + step]
        part = (r @ r).reshape(1)
        return 0.5 * float(allreduce_dev(part)[0]) + lam * sum(
            float(np.linalg.norm(sol["x"][k * gs:(k + 1) * gs])) for k in range(K))

    return ProxAffineProblem(prox, cons, name="group_lasso_%dx%d_device" % (m, n)), {
        "objective": objective, "solution_var": "x", "lam": lam, "sharded_args": {(0, 0): m}, "keep": a}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config", choices=["tv1d", "logreg", "hinge", "group_lasso", "lasso_dense", "lasso_sparse", "mv_lasso"])
    # --rows / --cols: spellings torchrun's own parser does not mistake for abbreviations of its options
    ap.add_argument("--m", "--rows", dest="m", type=int, default=0)
    ap.add_argument("--n", "--cols", dest="n", type=int, default=0)
    ap.add_argument("--density", type=float, default=1.0)
    ap.add_argument("--group", type=int, default=50)
    ap.add_argument("--K", type=int, default=100)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--kkt-mode", default="auto")
    ap.add_argument("--max-iterations", type=int, default=2000)
    ap.add_argument("--no-solve", action="store_true")
    ap.add_argument("--no-profile", action="store_true")
    ap.add_argument("--reuse-solver", action="store_true", help="time the iterations on the solver of the solve (one Init)")
    ap.add_argument("--device-gen", action="store_true", help="hinge: generate the dense data block on the GPU")
    a = ap.parse_args()
    # libraries print banners on fd 1 (NCCL's version line): keep the real stdout for the one JSON line
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    import torch
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world)

    from epsilon_b200 import problems
    from epsilon_b200.solver import Context, ProxADMMSolver

    def allreduce(v):
        if world == 1:
            return v
        t = torch.from_numpy(np.ascontiguousarray(v, dtype=np.float64).copy())
        dist.all_reduce(t)
        return t.numpy()

    def allreduce_dev(t):
        if world > 1:
            dist.all_reduce(t)
        return t

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    t0 = time.time()
    sharded_constraints = None
    sharded_args = None
    if a.config == "tv1d":
        prob, info = problems.tv_1d(a.n or 10_000_000, seed=0)
    elif a.config == "logreg" and (world > 1 or a.density < 1.0 and a.m >= 100_000):
        per = 8 // world
        prob, info = problems.logreg_l1_panels(a.m, a.n, a.density, panels=8, my_panels=range(rank * per, (rank + 1) * per),
                                               seed=0, allreduce=allreduce)
        sharded_constraints = info["sharded_constraints"] if world > 1 else None
    elif a.config == "logreg":
        prob, info = problems.logreg_l1(a.m, a.n, a.density, seed=0)
    elif a.config == "hinge" and (a.device_gen or world > 1):
        prob, info = hinge_l2_on_device(a.m, a.n, rank, world, 0, allreduce_dev)
        sharded_constraints = info["sharded_constraints"] if world > 1 else None
    elif a.config == "hinge":
        prob, info = problems.hinge_l2(a.m, a.n, a.density, seed=0)
    elif a.config == "group_lasso" and (a.device_gen or world > 1):
        prob, info = group_lasso_on_device(a.m, a.group, a.K, rank, world, 0, allreduce_dev)
        sharded_args = info["sharded_args"] if world > 1 else None
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

    ctx = Context(local_rank)
    if world > 1:
        uid = [Context.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.init_comm(world, rank, uid[0])
    params = {"kkt_mode": a.kkt_mode, "max_iterations": a.max_iterations}
    out = {"config": a.config, "problem": prob.name, "n_gpus": world, "data_generation_s": round(t_gen, 2)}
    kw = dict(ctx=ctx, sharded_constraints=sharded_constraints, sharded_args=sharded_args)

    if not a.no_solve:
        if world > 1:
            dist.barrier()
        t0 = time.time()
        s = ProxADMMSolver(prob, params, **kw)
        t_init = time.time() - t0
        x = s.solve()
        ctx.synchronize()
        t_all = max_over_ranks(time.time() - t0)
        st = s.status.as_dict()
        out["solve"] = {"state": st["state"], "num_iterations": st["num_iterations"],
                        "iterations_run": st["iterations_run"], "init_s": round(max_over_ranks(t_init), 4),
                        "loop_s": st["loop_ms"] / 1e3, "time_to_tol_s": round(t_all, 4),
                        "objective": info["objective"](x), "r_norm": st["r_norm"], "s_norm": st["s_norm"],
                        "cg_iterations": st["cg_iterations"], "cg_unconverged": st["cg_unconverged"],
                        "m": st["m"], "n": st["n"]}
        out["solver"] = s.describe().strip().split("\n")[:12]
        if not a.reuse_solver:
            s.close()

    if a.no_solve or not a.reuse_solver:
        s = ProxADMMSolver(prob, params, **kw)
    s.iterate(a.warmup)
    if world > 1:
        dist.barrier()
    ms = max_over_ranks(s.iterate(a.steps))
    out["iterations_per_s"] = a.steps / (ms / 1e3)
    out["ms_per_iteration"] = ms / a.steps
    out["cg_iterations_per_step"] = s.residuals()["cg_iterations"] / a.steps
    if "solver" not in out:
        out["solver"] = s.describe().strip().split("\n")[:12]
    s.close()

    if not a.no_profile:
        s = ProxADMMSolver(prob, dict(params, profile=1), **kw)
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
        s.close()
    out["hbm_peak_GBps"] = peak
    if world > 1:
        dist.barrier()
    ctx.close()
    if rank == 0:
        real_stdout.write(json.dumps(out) + "\n")
        real_stdout.flush()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
