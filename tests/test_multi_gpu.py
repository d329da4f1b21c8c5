"""Two ranks on two GPUs of one box (NCCL): the paths of activesplat_b200/dist.py that need real peers.
Skipped on a single-GPU box (run with `gpurun --gpus 2 -- python -m pytest tests/test_multi_gpu.py -m gpu`);
the host-side logic of the same functions is covered on CPU with gloo (tests/test_dist_cpu.py)."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers  # noqa: F401

pytestmark = pytest.mark.gpu


def _two_gpus():
    return torch.cuda.is_available() and torch.cuda.device_count() >= 2


def _peer_adam_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from activesplat_b200 import dist as asd
    from activesplat_b200.frame import FusedAdam, PARAM_ORDER
    P = 20_003                                      # not a multiple of 4: slabs cut through tensors and rows
    shapes = {"means3D": 3, "rgb_colors": 3, "unnorm_rotations": 4, "logit_opacities": 1, "log_scales": 3}
    lrs = {"means3D": 1e-4, "rgb_colors": 2.5e-3, "unnorm_rotations": 1e-3, "logit_opacities": 5e-2, "log_scales": 1e-3}
    g0 = torch.Generator(device=dev).manual_seed(1)            # same on every rank: the replicated map
    a = {k: torch.randn(P, w, generator=g0, device=dev) for k, w in shapes.items()}
    b = {k: v.clone() for k, v in a.items()}
    gr = torch.Generator(device=dev).manual_seed(100 + rank)   # per-rank gradients (this rank's keyframes)
    flat_a = torch.zeros(14 * P, device=dev)
    flat_b = torch.zeros(14 * P, device=dev)
    opt_a = FusedAdam(a, lrs)
    opt_b = asd.PeerShardedAdam(b, lrs, flat_b)
    for step in range(3):
        grad = torch.randn(14 * P, generator=gr, device=dev) * 10.0 ** (step - 1)
        flat_a.copy_(grad)
        flat_b.copy_(grad)
        asd.allreduce_gradients(flat_a)             # reference path: NCCL all-reduce, then the same Adam everywhere
        opt_a.step(flat_a)
        opt_b.step()                                # one kernel: reduce over peer memory + Adam + write to every rank
    torch.cuda.synchronize(dev)
    worst = max(float((a[k] - b[k]).abs().max() / lrs[k]) for k in PARAM_ORDER)   # in units of one Adam step
    sums = torch.stack([b[k].double().sum() for k in PARAM_ORDER])
    lo, hi = sums.clone(), sums.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    q.put((rank, worst, bool(torch.equal(lo, hi))))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(not _two_gpus(), reason="needs two GPUs")
def test_peer_sharded_adam_two_ranks_equals_allreduce_plus_adam():
    """as_adam_step_peer over CUDA-IPC peer mappings on two ranks: parameters equal the all-reduce + as_adam_step path
    (sums differ only in the order of two additions) and are bit-identical across ranks."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + (os.getpid() % 1000)
    procs = [ctx.Process(target=_peer_adam_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for rank, worst, same in out:
        assert same, f"rank {rank}: parameters differ between ranks"
        assert worst <= 2e-3, f"rank {rank}: {worst} Adam steps away from the all-reduce path"
