"""The row-sharded (multi-GPU) device path on hardware: CUDA-IPC peer buffers, the two-shot exchange with
flag-in-data packets inside CgFusedTailKernel, the one-shot protocol (EPS_PEER_LL=0), the NCCL fallback
(EPS_NO_PEER=1) and the generic sharded matvec (CSR kernels + ncclAllReduce per product).

One process per GPU (spawned here, world size 2 or more when the box has the GPUs); every rank holds a row
block of the data matrix.  Asserted: the per-epoch trace, the iteration count and the solution are identical
on every rank, and equal to the REFERENCE's fixture (tests/golden, rel 1e-7) or, off the fixtures, to the
pinned CPU oracle.  Skipped on a single-GPU box; profiles/ holds the log of a 2-GPU run."""
import json
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLDEN = os.path.join(HERE, "golden")
for p in (ROOT, GOLDEN):
    if p not in sys.path:
        sys.path.insert(0, p)

pytestmark = pytest.mark.gpu

TRACE_KEYS = ("state", "num_iterations", "r_norm", "s_norm", "epsilon_primal", "epsilon_dual")


def _gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


# case -> (generator args); the first one is a committed reference fixture
CASES = {
    "lasso_sparse_4000x400": dict(m=4000, n=400, density=0.02, seed=3),       # tests/golden fixture
    "lasso_sparse_odd_3001x301": dict(m=3001, n=301, density=0.03, seed=21),  # odd n, ragged row split
    "lasso_sparse_tiny_64x5": dict(m=64, n=5, density=0.5, seed=22),          # one tail CTA < ranks' slices
}

# exchange protocol -> environment of the worker processes
MODES = {
    "sentinel_two_shot": {"EPS_SPARSE_FORMAT": "sell"},                    # fused tail, 8-byte sentinel slots (default)
    "ll_two_shot": {"EPS_SPARSE_FORMAT": "sell", "EPS_PEER_SENTINEL": "0"},   # fused tail, 16-byte flag-in-data packets
    "ll_two_shot_v1": {"EPS_SPARSE_FORMAT": "sell", "EPS_TAIL_V1": "1"},   # first-generation tail (two grid syncs)
    "sentinel_grown": {"EPS_SPARSE_FORMAT": "sell", "EPS_PEER_CAPACITY": "64"},   # peer buffers grow in place at solver set-up
    "one_shot_flags": {"EPS_SPARSE_FORMAT": "sell", "EPS_PEER_LL": "0"},   # P2P stores + fence + per-CTA flags
    "nccl_fallback": {"EPS_SPARSE_FORMAT": "sell", "EPS_NO_PEER": "1"},    # ncclAllReduce per product
    "csr_nccl": {"EPS_SPARSE_FORMAT": "csr"},                              # generic block path, CSR kernels
}


def _worker(rank, world, port, env, case, out_dir):
    os.environ.update(env)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from epsilon_b200 import problems
    from epsilon_b200.solver import Context, ProxADMMSolver
    ctx = Context(rank)
    uid = [Context.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    ctx.init_comm(world, rank, uid[0])
    a = CASES[case]
    _, info = problems.lasso(a["m"], a["n"], density=a["density"], seed=a["seed"])
    prob, sharded = problems.lasso_row_shard(info, rank, world)
    res = {}
    # two solves on one context: the second reuses the peer buffers with advanced sequence numbers
    for rep in range(2):
        with ProxADMMSolver(prob, {"kkt_mode": "cg"}, ctx=ctx, sharded_args=sharded) as s:
            x = s.solve()
            st = s.status.as_dict()
            trace = np.array([[t[k] for k in TRACE_KEYS] for t in s.trace], dtype=np.float64)
            desc = s.describe()
        res["x%d" % rep] = x["x"]
        res["trace%d" % rep] = trace
        res["unconverged%d" % rep] = st["cg_unconverged"]
        res["cg%d" % rep] = st["cg_iterations"]
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), desc=np.array(desc), **res)
    dist.barrier()
    ctx.close()
    dist.destroy_process_group()


def _expected(case):
    import make_golden
    from oracle import prox_admm_np as onp
    from epsilon_b200 import problems
    a = CASES[case]
    index = json.load(open(os.path.join(GOLDEN, "index.json")))
    if case in index["solve"]:
        prob, info, params = make_golden.make_problem(case)
        if make_golden.problem_digest(prob) == index["solve"][case]["digest"]:
            g = np.load(os.path.join(GOLDEN, case + ".npz"))
            return g["trace"], g["x"], "reference fixture"
    prob, info = problems.lasso(a["m"], a["n"], density=a["density"], seed=a["seed"])
    o = onp.ProxADMM(prob).solve()
    trace = np.array([[t[k] for k in TRACE_KEYS] for t in o["trace"]], dtype=np.float64)
    return trace, o["x"]["x"], "CPU oracle"


@pytest.mark.timeout(900)
@pytest.mark.parametrize("mode", sorted(MODES))
@pytest.mark.parametrize("case", sorted(CASES))
def test_row_sharded_solve_matches_reference_on_every_rank(case, mode, tmp_path):
    world = min(_gpus(), int(os.environ.get("EPS_TEST_WORLD", "2")))
    if world < 2:
        pytest.skip("needs at least 2 GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(world, _free_port(), MODES[mode], case, str(tmp_path)), nprocs=world, join=True)
    r = [np.load(os.path.join(str(tmp_path), "rank%d.npz" % k)) for k in range(world)]
    desc = str(r[0]["desc"])
    if mode in ("sentinel_two_shot", "sentinel_grown", "ll_two_shot", "ll_two_shot_v1", "one_shot_flags"):
        assert "peer-allreduce" in desc and "fused-tail" in desc, desc
    if mode == "nccl_fallback":
        assert "sharded" in desc and "peer-allreduce" not in desc, desc
    gold_trace, gold_x, _ = _expected(case)
    for rep in range(2):
        # all ranks bit-identical: the sums are formed in rank order on every rank
        for k in range(1, world):
            assert np.array_equal(r[0]["trace%d" % rep], r[k]["trace%d" % rep]), (mode, rep, desc)
            assert np.array_equal(r[0]["x%d" % rep], r[k]["x%d" % rep]), (mode, rep)
        assert int(r[0]["unconverged%d" % rep]) == 0
        trace = r[0]["trace%d" % rep]
        assert trace.shape == gold_trace.shape, (trace.shape, gold_trace.shape, desc)
        assert np.array_equal(trace[:, :2], gold_trace[:, :2])           # states, iteration numbers
        np.testing.assert_allclose(trace[:, 2:], gold_trace[:, 2:], rtol=1e-7, atol=1e-12)
        x = r[0]["x%d" % rep]
        assert np.abs(x - gold_x).max() <= 1e-7 * max(1.0, np.abs(gold_x).max())
    # the second solve on the same context repeats the first bit for bit
    assert np.array_equal(r[0]["trace0"], r[0]["trace1"]) and np.array_equal(r[0]["x0"], r[0]["x1"])


# ---- row-sharded CONSTRAINTS (eps_problem_constraint_sharded): logistic regression and the SVM ------------------
SHARD_CASES = {
    # fixture name -> (generator, shard helper, params)
    "logreg_l1_2000x200": ("logreg_l1", "logreg_l1_row_shard", dict(m=2000, n=200, density=0.05, seed=4), {"max_iterations": 200}),
    "hinge_l2_sparse_5000x300": ("hinge_l2", "hinge_l2_row_shard", dict(m=5000, n=300, density=0.02, seed=10), {"max_iterations": 300}),
    "hinge_l2_1500x100": ("hinge_l2", "hinge_l2_row_shard", dict(m=1500, n=100, seed=5), {"max_iterations": 300}),   # dense data block
    "logreg_l1_sparse_60000x2000": ("logreg_l1", "logreg_l1_row_shard", dict(m=60000, n=2000, density=0.005, seed=14), {"max_iterations": 200}),
}
SHARD_MODES = {
    "gram": {"kkt_mode": "gram"},                                  # Gram of the row blocks all-reduced, factor replicated
    "cg_fused": {"kkt_mode": "cg", "_env": {"EPS_SPARSE_FORMAT": "sell"}},   # fused SELL passes + NVLink exchange per CG step
    "cg_nccl": {"kkt_mode": "cg", "_env": {"EPS_SPARSE_FORMAT": "csr"}},     # generic block path, ncclAllReduce per product
}


def _shard_worker(rank, world, port, case, mode, out_dir):
    gen, helper, args, params = SHARD_CASES[case]
    mp_ = dict(SHARD_MODES[mode])
    os.environ.update(mp_.pop("_env", {}))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from epsilon_b200 import problems
    from epsilon_b200.solver import Context, ProxADMMSolver
    ctx = Context(rank)
    uid = [Context.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    ctx.init_comm(world, rank, uid[0])
    a = dict(args)
    m, n = a.pop("m"), a.pop("n")
    _, info = getattr(problems, gen)(m, n, **a)
    prob, sharded = getattr(problems, helper)(info, rank, world)
    with ProxADMMSolver(prob, dict(params, **mp_), ctx=ctx, sharded_constraints=sharded) as s:
        x = s.solve()
        st = s.status.as_dict()
        trace = np.array([[t[k] for k in TRACE_KEYS] for t in s.trace], dtype=np.float64)
        desc = s.describe()
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), desc=np.array(desc), x=x["x"], t=x["t"], trace=trace,
             mn=np.array([st["m"], st["n"]]), unconverged=st["cg_unconverged"])
    dist.barrier()
    ctx.close()
    dist.destroy_process_group()


@pytest.mark.timeout(900)
@pytest.mark.parametrize("mode", sorted(SHARD_MODES))
@pytest.mark.parametrize("case", sorted(SHARD_CASES))
def test_row_sharded_constraints_match_the_reference_fixture(case, mode, tmp_path):
    """C3 / C5b shapes: the data matrix sits in the CONSTRAINT, so u, y_i and t are m-vectors and the rows of the
    constraint block are what is sharded.  Every rank must report the reference's trace (the single-process
    fixture), hold the replicated x bit-identically, and the shards of t must concatenate to the fixture's."""
    world = min(_gpus(), int(os.environ.get("EPS_TEST_WORLD", "2")))
    if world < 2:
        pytest.skip("needs at least 2 GPUs (gpurun --gpus 2)")
    if mode == "cg_fused" and "sparse" not in case and case != "logreg_l1_2000x200":
        pytest.skip("dense data block: no SELL path")
    import make_golden
    index = json.load(open(os.path.join(GOLDEN, "index.json")))
    prob, info, params = make_golden.make_problem(case)
    if make_golden.problem_digest(prob) != index["solve"][case]["digest"]:
        pytest.skip("generator drift on this host")
    gold = np.load(os.path.join(GOLDEN, case + ".npz"))
    import torch.multiprocessing as mp
    mp.spawn(_shard_worker, args=(world, _free_port(), case, mode, str(tmp_path)), nprocs=world, join=True)
    r = [np.load(os.path.join(str(tmp_path), "rank%d.npz" % k)) for k in range(world)]
    desc = str(r[0]["desc"])
    assert r[0]["mn"].tolist() == gold["mnN"][:2].tolist(), desc        # global A_.m(), A_.n()
    if mode == "cg_fused":
        assert "fused-matvec(constraint block)" in desc and "peer-allreduce" in desc and "fused-tail" in desc, desc
    for k in range(1, world):
        assert np.array_equal(r[0]["trace"], r[k]["trace"]), (mode, desc)
        assert np.array_equal(r[0]["x"], r[k]["x"])
    assert int(r[0]["unconverged"]) == 0
    trace = r[0]["trace"]
    assert trace.shape == gold["trace"].shape, (trace.shape, gold["trace"].shape, desc)
    assert np.array_equal(trace[:, :2], gold["trace"][:, :2])
    np.testing.assert_allclose(trace[:, 2:], gold["trace"][:, 2:], rtol=1e-7, atol=1e-12)
    assert np.abs(r[0]["x"] - gold["x"]).max() <= 1e-7 * max(1.0, np.abs(gold["x"]).max())
