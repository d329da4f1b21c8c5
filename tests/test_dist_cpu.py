"""CPU, world_size 2, gloo: the host-side sharding logic of activesplat_b200/dist.py
(view scoring gather and keyframe-batch gradient all-reduce) against the single-process result."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers  # noqa: F401


def _views_worker(rank, world, port, n_views, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from activesplat_b200 import dist as asd
    calls = []

    def render_fn(v, slot=0):  # a deterministic stand-in for the per-view opacity image
        calls.append(v)
        g = torch.Generator().manual_seed(1000 + v)
        return torch.rand(1, 15, 12, generator=g)

    full = asd.score_views(render_fn, n_views)
    q.put((rank, calls, full.clone()))
    dist.barrier()
    dist.destroy_process_group()


def _mapping_worker(rank, world, port, n_kf, q, weights=None):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from activesplat_b200 import dist as asd
    P = 50
    flat = torch.full((P, 14), float("nan"))
    params = torch.zeros(P, 14)

    def frame_grad_fn(k, accumulate):
        g = torch.Generator().manual_seed(77 + k)
        gk = torch.randn(P, 14, generator=g)
        if accumulate:
            flat.add_(gk)
        else:
            flat.copy_(gk)
        return gk.abs().sum()

    def opt_step(f):
        params.sub_(0.1 * f)

    asd.mapping_step_keyframe_batch(frame_grad_fn, n_kf, flat, opt_step, weights=weights)
    q.put((rank, params.clone()))
    dist.barrier()
    dist.destroy_process_group()


class _SgdStandIn:
    """step / step_range / n_flat of FusedAdam, as plain SGD on a CPU tensor (the chunking logic is what is tested)."""

    def __init__(self, params, n_flat):
        self.params, self.n_flat, self.calls, self.advanced = params, n_flat, [], 0

    def step(self, flat):
        self.step_range(flat, 0, self.n_flat, True)

    def step_range(self, flat, e0, e1, advance):
        self.calls.append((e0, e1))
        self.advanced += int(bool(advance))
        self.params[e0:e1].sub_(0.1 * flat[e0:e1])


def _chunked_worker(rank, world, port, n, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from activesplat_b200 import dist as asd
    flat = torch.randn(n + 7, generator=torch.Generator().manual_seed(500 + rank))   # tail beyond n_flat: untouched
    params = torch.zeros(n + 7)
    opt = _SgdStandIn(params, n)
    asd.allreduce_and_step(flat, opt, n_chunks=4)
    q.put((rank, params.clone(), opt.calls, opt.advanced))
    dist.barrier()
    dist.destroy_process_group()


def test_chunked_allreduce_and_step_equals_allreduce_then_step():
    for n in (4000, 4001, 10):
        res = _spawn(_chunked_worker, 2, n)
        total = sum(torch.randn(n + 7, generator=torch.Generator().manual_seed(500 + r)) for r in range(2))
        for r in range(2):
            _, params, calls, advanced = res[r]
            torch.testing.assert_close(params[:n], -0.1 * total[:n])
            assert float(params[n:].abs().max()) == 0.0 and advanced == 1
            assert calls[0][0] == 0 and calls[-1][1] == n and all(a[1] == b[0] for a, b in zip(calls, calls[1:]))
            assert all(c[0] % 4 == 0 for c in calls)


def _spawn(fn, world, *args):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    kw = args[-1] if args and isinstance(args[-1], dict) else {}
    args = args[:-1] if kw else args
    procs = [ctx.Process(target=fn, args=(r, world, port) + args + (q,), kwargs=kw) for r in range(world)]
    for p in procs:
        p.start()
    out = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return sorted(out, key=lambda t: t[0])


def test_view_scoring_shards_and_gathers():
    for n_views in (7, 2, 1):
        res = _spawn(_views_worker, 2, n_views)
        # each view rendered exactly once, round-robin
        assert res[0][1] == list(range(0, n_views, 2)) and res[1][1] == list(range(1, n_views, 2))
        expect = []
        for v in range(n_views):
            op = torch.rand(1, 15, 12, generator=torch.Generator().manual_seed(1000 + v))
            expect.append(torch.stack([(1 - op).sum(), op.mean()]))
        expect = torch.stack(expect)
        for r in range(2):
            torch.testing.assert_close(res[r][2], expect)


def test_keyframe_batch_mapping_allreduce_matches_sequential():
    for n_kf in (5, 1):
        res = _spawn(_mapping_worker, 2, n_kf)
        total = sum(torch.randn(50, 14, generator=torch.Generator().manual_seed(77 + k)) for k in range(n_kf))
        for r in range(2):
            torch.testing.assert_close(res[r][1], -0.1 * total, rtol=1e-6, atol=1e-6)


def test_weighted_deal_balances_and_gives_the_same_step():
    """deal_by_weight: every unit exactly once, loads within the heaviest unit of each other, identical on every rank;
    a keyframe batch dealt by weight yields the same update as the sequential sum."""
    from activesplat_b200 import dist as asd
    g = torch.Generator().manual_seed(3)
    for n, world in ((64, 8), (7, 2), (3, 4), (0, 2)):
        w = (torch.rand(n, generator=g) * 5 + 0.5).tolist()
        deal = asd.deal_by_weight(w, world)
        assert sorted(sum(deal, [])) == list(range(n)) and len(deal) == world
        loads = [sum(w[i] for i in d) for d in deal]
        if n:
            assert max(loads) - min(loads) <= max(w) + 1e-9
    w5 = [5.0, 1.0, 1.0, 1.0, 1.0]
    assert asd.deal_by_weight(w5, 2) == [[0], [1, 2, 3, 4]]
    res = _spawn(_mapping_worker, 2, 5, {"weights": w5})
    total = sum(torch.randn(50, 14, generator=torch.Generator().manual_seed(77 + k)) for k in range(5))
    for r in range(2):
        torch.testing.assert_close(res[r][1], -0.1 * total, rtol=1e-6, atol=1e-6)


def test_shard_indices_cover_exactly():
    from activesplat_b200 import dist as asd
    for n in (0, 1, 5, 256):
        for world in (1, 2, 4, 8):
            got = sorted(sum((asd.shard_indices(n, r, world) for r in range(world)), []))
            assert got == list(range(n))


def test_packed_cameras_equal_the_per_view_matrices():
    """dist._pack_cameras (one upload for a whole candidate-view sweep) against CameraSpec.matrices(), the per-view
    restatement of setup_camera (src/mapper/splatam/utils/recon_helpers.py:6-14): same fp32 values, bit for bit."""
    import numpy as np
    from activesplat_b200 import dist as asd, scenes
    views = scenes.candidate_views(7, seed=3)
    packed = asd._pack_cameras(views).numpy()
    assert packed.shape == (7, 54)
    for j, v in enumerate(views):
        view, proj, campos = v.matrices()
        assert np.array_equal(packed[j, 0:16].reshape(4, 4), view)
        assert np.array_equal(packed[j, 16:32].reshape(4, 4), proj)
        assert np.array_equal(packed[j, 32:48].reshape(4, 4), np.asarray(v.w2c, dtype=np.float32))
        assert np.array_equal(packed[j, 48:51], campos)
        assert (packed[j, 51:54] == 0).all()
