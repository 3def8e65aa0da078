"""Per-construct ingest timing under torchrun (row-sharded lasso), to see which solver construction is slow."""
import json, os, sys, time
import numpy as np, torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from epsilon_b200 import problems
from epsilon_b200.solver import Context, ProxADMMSolver
world, rank, lr = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world)
def allreduce(a):
    t = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64).copy()); dist.all_reduce(t); return t.numpy()
per = 8 // world
prob, info = problems.lasso_sparse_panels(1_000_000, 100_000, 1e-3, panels=8, my_panels=range(rank*per,(rank+1)*per), seed=0, allreduce=allreduce)
a = info["A"]; a.sort_indices()
pinned = [torch.from_numpy(v).pin_memory() for v in (a.data, a.indices, a.indptr)]
a.data, a.indices, a.indptr = (t.numpy() for t in pinned)
ctx = Context(lr)
uid = [Context.nccl_unique_id() if rank == 0 else None]; dist.broadcast_object_list(uid, src=0); ctx.init_comm(world, rank, uid[0])
for rep in range(6):
    ctx.synchronize(); dist.barrier()
    t0 = time.time()
    s = ProxADMMSolver(prob, {"cg_tol": 1e-12}, ctx=ctx, sharded_args=info["sharded_args"])
    ctx.synchronize(); t1 = time.time()
    s.solve(); ctx.synchronize(); t2 = time.time()
    s.close(); ctx.synchronize(); t3 = time.time()
    if rank == 0: print(json.dumps({"rep": rep, "construct_ms": round(1e3*(t1-t0),1), "solve_ms": round(1e3*(t2-t1),1), "close_ms": round(1e3*(t3-t2),2)}), flush=True)
dist.barrier(); ctx.close(); dist.destroy_process_group()
