"""torchrun script: sharded evaluation (world ranks, NCCL gather to rank 0) against the single-rank evaluation on rank 0."""
import ctypes, os, sys
import numpy as np
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
m = ge.package()
m.lib().mocob200_set_min_points_per_rank(int(os.environ.get("MIN_POINTS", "8")))  # shard the small fixtures too (same call on every rank)
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
def fresh_id():  # one NCCL unique id per communicator (= per evaluator)
    idbuf = torch.zeros(128, dtype=torch.uint8)
    if rank == 0:
        raw = ctypes.create_string_buffer(128)
        assert m.lib().mocob200_nccl_unique_id(raw) == 0
        idbuf = torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8).clone()
    idbuf = idbuf.cuda(); dist.broadcast(idbuf, 0)
    return bytes(idbuf.cpu().numpy().tobytes())
ok = True
for fixture, opts, fd in (("gait2d_tracking.mb20", None, 0), ("gait2d_prediction.mb20", "num_mesh_intervals=40", 1), ("pendulum3_implicit.mb20", None, 0),
                          ("pendulum_cb_both.mb20", None, 1), ("gait2d_prediction.mb20", "num_mesh_intervals=800", 1)):
    p = m.Problem.load(os.path.join(ROOT, "tests", "golden", fixture), opts)
    x = p.make_iterate(seed=7, spread=0.2)
    ev = m.Evaluator(p, device=local, rank=rank, world=world, nccl_id=fresh_id())
    ev.set_fd(fd)
    if rank == 0:
        print(fixture, opts, "working ranks", ev.working_ranks(), "of", world, "warp kernels" if ev.uses_warp_kernels() else "thread kernels")
    g, v, gr, f = ev.eval_g(x), ev.eval_jac_g(x), ev.eval_grad_f(x), ev.eval_f(x)
    if rank == 0:
        ev1 = m.Evaluator(p, device=local)
        ev1.set_fd(fd)
        g1, v1, gr1, f1 = ev1.eval_g(x), ev1.eval_jac_g(x), ev1.eval_grad_f(x), ev1.eval_f(x)
        same = np.array_equal(g, g1) and np.array_equal(v, v1, equal_nan=True) and np.array_equal(gr, gr1, equal_nan=True) and f == f1
        print(fixture, "world", world, "sharded == single-rank:", same, "max|dv|", np.nanmax(np.abs(v - v1)))
        ok = ok and same
    dist.barrier()
if rank == 0:
    print("MULTIRANK_OK" if ok else "MULTIRANK_MISMATCH")
dist.destroy_process_group()
