"""Host-side logic of the row-sharded (N > 1) path on CPU: world size 2 over gloo.

What the device library does across ranks (csrc/kkt.cu, comm.cu) is: every rank holds a row
block of the data matrix A and of b, all n-vectors are replicated, and the only exchange is
sum-all-reduce of A_r'(A_r p) per CG step (plus A_r'b_r and scalars at Init).  Here the same
partition is driven with the panel generator bench.py uses and a numpy emulation of that
exchange, and checked against the single-process CPU oracle on the stacked problem."""
import os
import socket
import sys

import numpy as np
import pytest
import scipy.sparse as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

M, N, DENS, PANELS = 4000, 300, 0.02, 4


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _sharded_admm(a_loc, b_loc, lam, n, allreduce, iters):
    """Lasso ADMM in the compiled form (SUM_SQUARE(A xs - b) + NORM_1[lam](x), xs - x = 0;
    prox_admm.cc:134-147) with the SUM_SQUARE prox solved by CG on (I + 2 A'A), the A'A product
    summed over ranks -- the exchange pattern of KktSolver::MatVec with h_sharded."""
    atb = allreduce(np.asarray(a_loc.T @ b_loc).ravel())
    u = np.zeros(n)
    y0 = np.zeros(n)
    y1 = np.zeros(n)
    xs = np.zeros(n)

    def matvec(p):
        return p + 2.0 * allreduce(np.asarray(a_loc.T @ (a_loc @ p)).ravel())

    for _ in range(iters):
        u = u - y0 - y1
        u = u + y0
        rhs = u + 2.0 * atb            # A'v - 2a H'g with A = I, g = -b, a = 1
        r = rhs - matvec(xs)
        p = r.copy()
        rr = r @ r
        bn = np.sqrt(rhs @ rhs)
        k = 0
        while np.sqrt(rr) > 1e-12 * bn and k < 500:
            w = matvec(p)
            al = rr / (p @ w)
            xs = xs + al * p
            r = r - al * w
            rr_new = r @ r
            p = r + (rr_new / rr) * p
            rr = rr_new
            k += 1
        y0 = xs.copy()
        u = u - y0
        u = u + y1
        v = -u
        x = np.sign(v) * np.maximum(np.abs(v) - lam, 0.0)
        y1 = -x
        u = u - y1
    return x, xs


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from epsilon_b200 import problems

    def allreduce(a):
        t = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64).copy())
        dist.all_reduce(t)
        return t.numpy()

    per = PANELS // world
    prob, info = problems.lasso_sparse_panels(M, N, DENS, panels=PANELS,
                                              my_panels=range(rank * per, (rank + 1) * per),
                                              seed=3, allreduce=allreduce)
    a, b = info["A"], info["b"]
    assert a.shape == (M // world, N)
    assert info["sharded_args"] == {(0, 0): M}
    x, xs = _sharded_admm(a, b, info["lam"], N, allreduce, iters=21)
    obj = info["objective"]({"x": x})
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), x=x, lam=info["lam"], obj=obj,
             indptr=a.indptr, indices=a.indices, data=a.data, b=b)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_row_sharded_lasso_world2_matches_single_process(tmp_path):
    import torch.multiprocessing as mp
    from epsilon_b200 import problems
    from oracle import prox_admm_np as onp

    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    r = [np.load(os.path.join(str(tmp_path), "rank%d.npz" % k)) for k in range(world)]

    # the global problem does not depend on how the panels are distributed
    prob, info = problems.lasso_sparse_panels(M, N, DENS, panels=PANELS, seed=3)
    shards = [sp.csc_matrix((q["data"], q["indices"], q["indptr"]), shape=(M // world, N)) for q in r]
    stacked = sp.vstack(shards).tocsc()
    assert abs(stacked - info["A"]).max() <= 1e-15
    assert np.array_equal(np.concatenate([q["b"] for q in r]), info["b"]) or \
        np.abs(np.concatenate([q["b"] for q in r]) - info["b"]).max() <= 1e-14
    assert float(r[0]["lam"]) == pytest.approx(info["lam"], rel=1e-13)
    # replicated n-space state is identical on both ranks
    assert np.array_equal(r[0]["x"], r[1]["x"])
    assert float(r[0]["obj"]) == float(r[1]["obj"])

    # and equals the single-process oracle after the same number of iterations
    o = onp.ProxADMM(prob, {"max_iterations": 21})
    for _ in range(21):
        o.iteration()
    x_ref = o.solution()["x"]
    assert np.abs(r[0]["x"] - x_ref).max() <= 1e-9 * max(1.0, np.abs(x_ref).max())
    assert float(r[0]["obj"]) == pytest.approx(info["objective"]({"x": x_ref}), rel=1e-9)
