"""Multi-GPU sharding of the hot path over one 8xB200 box (one process per GPU, torch.distributed/NCCL).

Two axes shard naturally; the reference itself is single-GPU (no distributed code):
  * candidate-view / keyframe RENDERS are independent given the (replicated) map: views are dealt
    round-robin to ranks, no collective on the data path, one tiny all_gather of the per-view scores
    (the GetOpacity sweep of src/visualizer/visualizer.py:991-1001 + src/mapper/splatam/__init__.py:698-759);
  * MAPPING over a keyframe batch: every rank accumulates the flat (14 P floats) parameter gradient of its
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


def deal_by_weight(weights, world):
    """Deal independent units of known, unequal cost to `world` ranks: heaviest first, each to the rank with the
    least work so far (longest-processing-time-first).  Deterministic, so every rank computes the same deal.
    Returns one ascending index list per rank.  Keyframes and candidate views of one map differ several-fold in
    Gaussian-tile instances; dealt round-robin, the step waits for whichever rank drew the heavy ones."""
    order = sorted(range(len(weights)), key=lambda i: (-float(weights[i]), i))
    loads = [0.0] * world
    out = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda q: (loads[q], q))
        out[r].append(i)
        loads[r] += float(weights[i])
    return [sorted(o) for o in out]


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
    """Sum the flat per-Gaussian gradient buffer (14 P floats) across ranks (one collective per mapping step)."""
    _, world = world_info(group)
    if world > 1:
        dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM, group=group)
    return flat_grad


def allreduce_and_step(flat_grad, optimizer, n_chunks=1, group=None):
    """Sum the flat gradient across ranks and apply the optimizer: the exchange + update of a mapping step.
    n_chunks > 1 pipelines the two: the buffer is all-reduced in pieces issued back to back on NCCL's stream and each
    piece is updated (FusedAdam.step_range) as soon as it has arrived, while the next ones are still on the wire.  Same
    result (the update is elementwise).  Measured on 2 B200s with a 56 MB gradient (1M Gaussians): one piece 0.874 ms
    per step, two 0.901, four 0.958 -- every extra NCCL launch costs more than the overlap returns, also inside a CUDA
    graph -- so one piece is the default; the option is kept for larger maps and slower links."""
    _, world = world_info(group)
    flat = flat_grad.reshape(-1)
    if world == 1 or n_chunks <= 1:
        if world > 1:
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        optimizer.step(flat)
        return
    n = optimizer.n_flat
    per = (-(-n // n_chunks) + 3) // 4 * 4
    bounds = [(b, min(n, b + per)) for b in range(0, n, per)]
    works = [dist.all_reduce(flat[b0:b1], op=dist.ReduceOp.SUM, group=group, async_op=True) for b0, b1 in bounds]
    for i, ((b0, b1), w) in enumerate(zip(bounds, works)):
        w.wait()   # stream-ordered: the current stream waits for this chunk, the host does not block
        optimizer.step_range(flat, b0, b1, advance=(i == len(bounds) - 1))


def mapping_step_keyframe_batch(frame_grad_fn, n_keyframes, flat_grad, optimizer_step, group=None, weights=None):
    """One mapping step over a keyframe batch split across ranks.
    frame_grad_fn(k, accumulate) must add keyframe k's gradient into `flat_grad` (overwriting when
    accumulate is False) and return its loss; optimizer_step(flat_grad) applies the update.
    weights: per-keyframe cost (e.g. the Gaussian-tile instances of the previous step) for a balanced deal
    (deal_by_weight); without it keyframes go round-robin.
    Equivalent single-GPU computation: accumulate all n_keyframes sequentially, then one step."""
    rank, world = world_info(group)
    mine = shard_indices(n_keyframes, rank, world) if weights is None else deal_by_weight(weights, world)[rank]
    loss = None
    for j, k in enumerate(mine):
        l = frame_grad_fn(k, j > 0)
        loss = l if loss is None else loss + l
    if not mine:
        flat_grad.zero_()
    allreduce_gradients(flat_grad, group)
    optimizer_step(flat_grad)
    return loss


# ---- candidate-view sweep with low host cost per view ---------------------------------------------------------
_VIEW_WORKSPACES = {}


def _pack_cameras(views):
    """setup_camera's matrices (recon_helpers.py:6-14, fp32 like the reference) for all views at once, packed
    [n, 54] = viewmatrix^T (16) | full projection^T (16) | w2c row-major (16) | camera centre (3) | zeros (3)."""
    import numpy as np
    w2c = torch.as_tensor(np.stack([np.asarray(v.w2c) for v in views])).float()
    n = w2c.shape[0]
    cam_center = torch.inverse(w2c)[:, :3, 3]
    view = w2c.transpose(1, 2)
    proj = torch.zeros((n, 4, 4))
    for i, v in enumerate(views):
        w, h, near, far = v.width, v.height, v.near, v.far
        proj[i] = torch.tensor([[2 * v.fx / w, 0.0, -(w - 2 * v.cx) / w, 0.0],
                                [0.0, 2 * v.fy / h, -(h - 2 * v.cy) / h, 0.0],
                                [0.0, 0.0, far / (far - near), -(far * near) / (far - near)],
                                [0.0, 0.0, 1.0, 0.0]])
    full = view.bmm(proj.transpose(1, 2))
    return torch.cat([view.reshape(n, 16), full.reshape(n, 16), w2c.reshape(n, 16), cam_center, torch.zeros(n, 3)], 1)


def score_candidate_views(params, views, group=None, n_streams=32, scale_modifier=1.0, return_opacity=False,
                          return_depth=False, shard=True):
    """The GetOpacity sweep (src/mapper/splatam/__init__.py:716-738 per view) over `views` (CameraSpec-like, all the
    same size), sharded over the ranks of `group`; every rank receives [n_views, 2] = (sum of invisibility, mean
    opacity).  Same renders as splatam.render_view, organised so that the host does little per view: all camera
    matrices are built and uploaded in one go, each view is one fused forward on one of `n_streams` streams (a
    low-resolution view is latency-bound on a few crowded tiles; in flight together they fill the GPU), and the
    scores are reduced once for the whole shard.
    return_opacity / return_depth: also return this rank's opacity / median-depth images [n_local,H,W].
    shard=False: render every view on this rank (the caller has already split the work) and skip the gather."""
    from . import splatam
    from .frame import FrameWorkspace, identity_pose
    rank, world = world_info(group) if shard else (0, 1)
    mine = shard_indices(len(views), rank, world)
    dev = params["means3D"].device
    P = int(params["means3D"].shape[0])
    W, H = int(views[0].width), int(views[0].height)
    n_streams = max(1, min(n_streams, len(mine)))
    key = (W, H, dev.index, n_streams)   # not P: the workspaces follow a growing map (FrameWorkspace.resize)
    if key not in _VIEW_WORKSPACES:
        _VIEW_WORKSPACES[key] = ([FrameWorkspace(P, W, H, dev) for _ in range(n_streams)],
                                 [torch.cuda.Stream(device=dev) for _ in range(n_streams)])
    wss, streams = _VIEW_WORKSPACES[key]
    rot, tr = identity_pose(dev)
    p = {k: params[k].detach() for k in ("means3D", "rgb_colors", "unnorm_rotations", "logit_opacities", "log_scales")}
    p["cam_unnorm_rots"], p["cam_trans"] = rot, tr
    white = splatam._white_bg(dev)
    ops = torch.empty((len(mine), H, W), dtype=torch.float32, device=dev)
    stats = torch.zeros((len(mine), 4), dtype=torch.int32, device=dev)
    depths = torch.empty((len(mine), H, W), dtype=torch.float32, device=dev) if return_depth else None
    if mine:
        d = _pack_cameras([views[v] for v in mine]).pin_memory().to(dev, non_blocking=True)
        cur = torch.cuda.current_stream(dev)
        for st in streams:
            st.wait_stream(cur)
        with torch.no_grad():
            for j, v in enumerate(mine):
                c = views[v]
                row = d[j]
                cam = splatam.Camera(image_height=H, image_width=W, tanfovx=W / (2 * c.fx), tanfovy=H / (2 * c.fy),
                                     bg=row[51:54], scale_modifier=scale_modifier, viewmatrix=row[0:16].view(1, 4, 4),
                                     projmatrix=row[16:32].view(1, 4, 4), sh_degree=0, campos=row[48:51],
                                     prefiltered=False, debug=False)
                s = j % n_streams
                with torch.cuda.stream(streams[s]):
                    wss[s].forward(p, cam, 0, sil_w2c=row[32:48].view(4, 4), bg_rgb=white, forward_only=True)
                    ops[j].copy_(wss[s].opacity[0], non_blocking=True)
                    if return_depth:
                        depths[j].copy_(wss[s].depth[0], non_blocking=True)
                    stats[j].copy_(wss[s].status, non_blocking=True)
        for st in streams:
            cur.wait_stream(st)
        # a view whose instances did not fit its workspace's binning buffer rendered as background: one look at all
        # status words after the sweep, and those views are rendered again with the buffer grown
        host = stats.cpu()
        for j in torch.nonzero(host[:, 1]).flatten().tolist():
            c, row, ws = views[mine[j]], d[j], wss[j % n_streams]
            cam = splatam.Camera(image_height=H, image_width=W, tanfovx=W / (2 * c.fx), tanfovy=H / (2 * c.fy),
                                 bg=row[51:54], scale_modifier=scale_modifier, viewmatrix=row[0:16].view(1, 4, 4),
                                 projmatrix=row[16:32].view(1, 4, 4), sh_degree=0, campos=row[48:51],
                                 prefiltered=False, debug=False)
            ws.ensure_capacity(int(host[j, 0]) + int(host[j, 0]) // 4)
            with torch.no_grad():
                ws.forward(p, cam, 0, sil_w2c=row[32:48].view(4, 4), bg_rgb=white, forward_only=True)
            if ws.last_status()[1]:
                raise RuntimeError("activesplat_b200: candidate view still overflows its binning buffer after growing it")
            ops[j].copy_(ws.opacity[0])
            if return_depth:
                depths[j].copy_(ws.depth[0])
    local = torch.stack([(1.0 - ops).sum((1, 2)), ops.mean((1, 2))], 1) if mine else torch.zeros((0, 2), device=dev)
    scores = gather_scores(local, len(views), rank, world, group)
    out = (scores,) + ((ops,) if return_opacity else ()) + ((depths,) if return_depth else ())
    return out if len(out) > 1 else scores


# ---- Adam fused with the gradient exchange over peer memory ------------------------------------------------------
class PeerShardedAdam:
    """Drop-in for `allreduce_gradients(flat) ; FusedAdam.step(flat)` when mapping over a keyframe batch on the GPUs of
    one box: each rank owns a contiguous slab of the flat buffer, sums that slab of every rank's gradient straight out of the
    peers' memory, updates it, and writes the new parameters into every rank's tensors (as_adam_step_peer).  Optimizer
    state exists only for the owned slab (1/world of FusedAdam's).  Results equal the all-reduce path up to the order of
    the float additions.

    The parameter tensors and the flat gradient buffer are exported once with CUDA IPC and opened by every peer on ITS
    device (as_ipc_open), so they must stay alive and in place; call `step()` on every rank."""

    def __init__(self, params, lrs, flat_grad, group=None, betas=(0.9, 0.999), eps=1e-15):
        import ctypes
        from . import capi
        from .frame import PARAM_ORDER
        self.rank, self.world = world_info(group)
        if self.world > 8:
            raise ValueError("PeerShardedAdam supports up to 8 ranks (one NVSwitch box)")
        self.group, self.params, self.flat_grad = group, params, flat_grad
        self.betas, self.eps, self.t = betas, eps, 0
        self.lrs = (ctypes.c_float * 5)(*[float(lrs[k]) for k in PARAM_ORDER])
        self.P = int(params["means3D"].shape[0])
        dev = params["means3D"].device
        n_flat = (12 if params["log_scales"].shape[1] == 1 else 14) * self.P   # elements in use (segment layout)
        per = -(-n_flat // self.world)
        per = (per + 3) // 4 * 4  # slabs start on 16-byte boundaries of the flat buffers
        self.g_begin, self.g_end = min(n_flat, per * self.rank), min(n_flat, per * (self.rank + 1))
        n_own = self.g_end - self.g_begin
        self.exp_avg = torch.zeros(n_own, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(n_own, dtype=torch.float32, device=dev)
        self._flag = torch.zeros(1, dtype=torch.float32, device=dev)
        mine = [params[k] for k in PARAM_ORDER] + [flat_grad]
        for t in mine:
            assert t.is_cuda and t.is_contiguous() and t.dtype == torch.float32
        own_ptrs = [t.data_ptr() for t in mine]
        self._opened = {}
        if self.world == 1:
            ptrs = [own_ptrs]
        else:
            # export: CUDA IPC handle of the allocation each tensor lives in + the tensor's byte offset inside it
            exports = []
            for t in mine:
                st = t.untyped_storage()
                _dev, handle, _size, off_bytes = st._share_cuda_()[:4]
                handle = bytes(handle)
                if len(handle) == 66:  # torch >= 2.3: version byte + allocation-kind byte + the 64-byte cudaIpcMemHandle_t
                    if handle[1:2] != b"c":
                        raise RuntimeError("PeerShardedAdam needs cudaMalloc-backed tensors (expandable segments are "
                                           "exported through a different mechanism)")
                    handle = handle[2:]
                if len(handle) != 64:
                    raise RuntimeError(f"unexpected CUDA IPC handle size {len(handle)}")
                exports.append((handle, int(off_bytes) + t.storage_offset() * t.element_size()))
            gathered = [None] * self.world
            dist.all_gather_object(gathered, exports, group=group)
            ptrs = []
            with torch.cuda.device(dev):  # open on THIS device: its kernels dereference the mapping (lazy peer access)
                for r in range(self.world):
                    if r == self.rank:
                        ptrs.append(own_ptrs)
                        continue
                    row = []
                    for handle, off in gathered[r]:
                        if handle not in self._opened:
                            base = ctypes.c_void_p()
                            capi.check(capi.lib().as_ipc_open(handle, ctypes.byref(base)))
                            self._opened[handle] = base.value
                        row.append(self._opened[handle] + off)
                    ptrs.append(row)
            torch.cuda.synchronize(dev)
        self._pp = (ctypes.c_void_p * (5 * self.world))(*[ptrs[r][t] for r in range(self.world) for t in range(5)])
        self._gp = (ctypes.c_void_p * self.world)(*[ptrs[r][5] for r in range(self.world)])

    def _barrier(self):
        if self.world > 1:  # stream-ordered: completes on a rank only after every rank's earlier work on its stream
            dist.all_reduce(self._flag, group=self.group)

    @torch.no_grad()
    def step(self):
        from . import capi
        self.t += 1
        p = self.params
        dev = p["means3D"].device
        self._barrier()   # every rank's gradient buffer is complete
        with torch.cuda.device(dev):
            capi.check(capi.lib().as_adam_step_peer(
                self.P, 1 if p["log_scales"].shape[1] == 1 else 0, self.world, self.rank, self.g_begin, self.g_end,
                self._pp, self._gp, self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(), self.lrs, self.betas[0],
                self.betas[1], self.eps, self.t, torch.cuda.current_stream(dev).cuda_stream))
        self._barrier()   # every rank's slab has landed in everybody's parameters
