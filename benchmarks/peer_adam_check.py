"""Multi-GPU check of dist.PeerShardedAdam (run under torchrun, one process per GPU):
every rank draws its own gradient; the peer-memory reduce+Adam must leave all ranks with the parameters that
all-reduce + FusedAdam produce (up to float summation order), and identical across ranks.  Also times both.

  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 benchmarks/peer_adam_check.py
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from activesplat_b200 import dist as asd  # noqa: E402
from activesplat_b200.frame import FusedAdam, PARAM_ORDER  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=dev)
P = int(os.environ.get("PEER_P", "3000000"))
shapes = {"means3D": 3, "rgb_colors": 3, "unnorm_rotations": 4, "logit_opacities": 1, "log_scales": 3}
lrs = {"means3D": 1e-4, "rgb_colors": 2.5e-3, "unnorm_rotations": 1e-3, "logit_opacities": 5e-2, "log_scales": 1e-3}
g0 = torch.Generator(device=dev).manual_seed(1)           # same on every rank: the replicated map
a = {k: torch.randn(P, w, generator=g0, device=dev) for k, w in shapes.items()}
b = {k: v.clone() for k, v in a.items()}
gr = torch.Generator(device=dev).manual_seed(100 + rank)  # per-rank gradients
flat_a = torch.zeros((P, 14), device=dev)
flat_b = torch.zeros((P, 14), device=dev)
opt_a = FusedAdam(a, lrs)
opt_b = asd.PeerShardedAdam(b, lrs, flat_b)
res = {"world": world, "P": P}
for step in range(3):
    grad = torch.randn(P, 14, generator=gr, device=dev)
    flat_a.copy_(grad)
    flat_b.copy_(grad)
    asd.allreduce_gradients(flat_a)
    opt_a.step(flat_a)
    opt_b.step()
torch.cuda.synchronize(dev)
worst = 0.0
for k in PARAM_ORDER:
    d = (a[k] - b[k]).abs().max() / (lrs[k] + 1e-30)     # in units of one Adam step of that tensor
    worst = max(worst, float(d))
    chk = b[k].double().sum().reshape(1)
    lo, hi = chk.clone(), chk.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    assert float(lo) == float(hi), f"ranks disagree on {k}"
res["max_diff_in_adam_steps"] = worst


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(dev)
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize(dev)
    t = torch.tensor([e0.elapsed_time(e1) / n], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


res["allreduce_plus_adam_ms"] = timeit(lambda: (asd.allreduce_gradients(flat_a), opt_a.step(flat_a)))
res["peer_reduce_adam_ms"] = timeit(lambda: opt_b.step())
if rank == 0:
    print(json.dumps(res), flush=True)
assert worst < 0.05, worst   # Adam with eps 1e-15 is sign-like: a different summation order moves an element by << 1 step
dist.barrier()
dist.destroy_process_group()
