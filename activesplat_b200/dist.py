"""Multi-GPU sharding of the hot path over one 8xB200 box (one process per GPU, torch.distributed/NCCL).

Two axes shard naturally; the reference itself is single-GPU (no distributed code):
  * candidate-view / keyframe RENDERS are independent given the (replicated) map: views are dealt
    round-robin to ranks, no collective on the data path, one tiny all_gather of the per-view scores
    (the GetOpacity sweep of src/visualizer/visualizer.py:991-1001 + src/mapper/splatam/__init__.py:698-759);
  * MAPPING over a keyframe batch: every rank accumulates the flat [P,14] parameter gradient of its
    keyframes, one all_reduce(SUM) over NVLink, then the identical fused Adam step on every rank.
The host-side logic here is backend-agnostic (tested with gloo on CPU); the renders are injected.
"""
import torch
import torch.distributed as dist


def world_info(group=None):
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def shard_indices(n, rank, world):
    """Round-robin deal of n independent units (views, keyframes) to `world` ranks."""
    return list(range(rank, n, world))


def gather_scores(local_scores, n_total, rank, world, group=None):
    """local_scores: [n_local, C] for units shard_indices(n_total, rank, world), in that order.
    Returns the full [n_total, C] table on every rank (rows in unit order)."""
    C = local_scores.shape[1] if local_scores.dim() == 2 else 1
    local_scores = local_scores.reshape(-1, C)
    if world == 1:
        return local_scores
    n_max = (n_total + world - 1) // world
    pad = torch.zeros((n_max, C), dtype=local_scores.dtype, device=local_scores.device)
    pad[: local_scores.shape[0]] = local_scores
    out = torch.empty((world, n_max, C), dtype=local_scores.dtype, device=local_scores.device)
    dist.all_gather_into_tensor(out.view(world * n_max, C), pad, group=group)
    full = torch.empty((n_total, C), dtype=local_scores.dtype, device=local_scores.device)
    for r in range(world):
        idx = shard_indices(n_total, r, world)
        full[idx] = out[r, : len(idx)]
    return full


def score_views(render_fn, n_views, device="cpu", group=None, n_streams=1):
    """GetOpacity scoring, sharded.  render_fn(view_index, slot) -> opacity image [1,H,W] (or any tensor whose
    `1 - x` summed is the view's invisibility, src/mapper/splatam/__init__.py:738).  Every rank renders
    its shard and all ranks receive [n_views, 2] = (sum of invisibility, mean opacity).

    n_streams > 1 (CUDA only): views are issued round-robin on that many streams, `slot` telling render_fn
    which of its per-stream workspaces to use.  A low-resolution candidate view has few tiles and ends in a
    long tail on a handful of SMs; with several views in flight the tails overlap instead of idling the GPU.
    """
    rank, world = world_info(group)
    mine = shard_indices(n_views, rank, world)
    local = torch.zeros((len(mine), 2), dtype=torch.float32, device=device)
    use_streams = n_streams > 1 and torch.device(device).type == "cuda"
    if use_streams:
        cur = torch.cuda.current_stream(device)
        streams = [torch.cuda.Stream(device=device) for _ in range(n_streams)]
        for st in streams:
            st.wait_stream(cur)
    for j, v in enumerate(mine):
        slot = j % n_streams if use_streams else 0
        if use_streams:
            with torch.cuda.stream(streams[slot]):
                op = render_fn(v, slot)
                local[j, 0] = (1.0 - op).sum()
                local[j, 1] = op.mean()
        else:
            op = render_fn(v, slot)
            local[j, 0] = (1.0 - op).sum()
            local[j, 1] = op.mean()
    if use_streams:
        for st in streams:
            cur.wait_stream(st)
    return gather_scores(local, n_views, rank, world, group)


def allreduce_gradients(flat_grad, group=None):
    """Sum the flat [P,14] per-Gaussian gradient buffer across ranks (one collective per mapping step)."""
    _, world = world_info(group)
    if world > 1:
        dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM, group=group)
    return flat_grad


def mapping_step_keyframe_batch(frame_grad_fn, n_keyframes, flat_grad, optimizer_step, group=None):
    """One mapping step over a keyframe batch split across ranks.
    frame_grad_fn(k, accumulate) must add keyframe k's gradient into `flat_grad` (overwriting when
    accumulate is False) and return its loss; optimizer_step(flat_grad) applies the update.
    Equivalent single-GPU computation: accumulate all n_keyframes sequentially, then one step."""
    rank, world = world_info(group)
    mine = shard_indices(n_keyframes, rank, world)
    loss = None
    for j, k in enumerate(mine):
        l = frame_grad_fn(k, j > 0)
        loss = l if loss is None else loss + l
    if not mine:
        flat_grad.zero_()
    allreduce_gradients(flat_grad, group)
    optimizer_step(flat_grad)
    return loss
