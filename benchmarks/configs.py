#!/usr/bin/env python
"""The five BASELINE.json configs, ours next to the reference pipeline on the same GPU.

  python benchmarks/configs.py [--configs 1,3,4] [--out gpurun_out/configs.json]
  torchrun --nproc-per-node N benchmarks/configs.py --configs 4,5      (multi-GPU legs)

"reference pipeline" = the reference's torch code restated (oracle/splatam_oracle.py) driving the
UNMODIFIED reference CUDA rasterizer built for sm_100a (oracle/_ref): two raster passes per frame,
torch-eager glue, torch.optim.Adam -- what ActiveSplat runs today.  It is a measured baseline, not a
product path.  bench.py is the contract benchmark (cfg2); this script fills in the other rows.
"""
import argparse
import importlib.util
import json
import math
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from activesplat_b200 import capi, dist as asd  # noqa: E402
from activesplat_b200 import scenes, splatam  # noqa: E402
from activesplat_b200.frame import PARAM_ORDER, FusedAdam, workspace_for  # noqa: E402

LW = dict(im=0.5, depth=1.0)
MAP_LRS = dict(means3D=1e-4, rgb_colors=2.5e-3, unnorm_rotations=1e-3, logit_opacities=5e-2, log_scales=1e-3,
               cam_unnorm_rots=0.0, cam_trans=0.0)  # config/splatam/online_habitat_sim.py:61-69


def load_ref_renderer(contiguous_inputs=False):
    """GaussianRasterizer-shaped class bound to the reference's _C (None if the build is absent).

    contiguous_inputs=True makes scales/rotations contiguous before the call.  The stock reference reads them
    with .contiguous() in its forward (dgr/rasterize_points.cu:102,104) but with a bare data_ptr() in its backward
    (:173,:175); SplaTAM's anisotropic path hands it rotations = F.normalize(quat_mult(...)) whose memory is
    column-major (quat_mult ends in torch.stack([...]).T, slam_helpers.py:28), so the stock backward computes
    dL/dscale and dL/drot from scrambled quaternions.  Our kernels (and the oracle) use the values the forward used."""
    path = os.path.join(ROOT, "oracle", "_ref", "ref_dgr_C.so")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location("ref_dgr_C", path)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    from activesplat_b200.rasterizer import make_rasterize_function
    fn = make_rasterize_function(ref)
    e = torch.Tensor([])

    class RefRenderer:
        def __init__(self, raster_settings):
            self.rs = raster_settings

        def __call__(self, means3D, means2D, opacities, colors_precomp, scales, rotations):
            if contiguous_inputs:
                scales, rotations = scales.contiguous(), rotations.contiguous()
            return fn.apply(means3D, means2D, e, colors_precomp, opacities, scales, rotations, e, self.rs)

    return RefRenderer


def timeit(fn, steps, warmup, dev):
    for i in range(warmup):
        fn(i)
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    torch.cuda.synchronize(dev)
    return e0.elapsed_time(e1) / steps


def gt_frames(params_cpu, cams, dev, seed):
    gtp = {k: v.to(dev) for k, v in scenes.perturbed(params_cpu, seed=seed).items()}
    out = []
    with torch.no_grad():
        for c in cams:
            im, depth, opacity, _ = splatam.render_view(gtp, c.w2c, c.k(), c.width, c.height)
            # render_view uses a white background; the GT colour is what it is -- only depth validity matters
            out.append(torch.cat([im, torch.where(opacity > 0.5, depth, torch.zeros_like(depth))], 0).clone())
    return out


# ---------------------------------------------------------------------------------------------------
def cfg1(dev, RefRenderer, args):
    """100k Gaussians, one 640x480 RGB-D+sil bundle fwd+bwd at the rasterizer boundary."""
    from activesplat_b200.rasterizer import GaussianRasterizer
    W, H, P = 640, 480, 100_000
    params, cam = scenes.s_frustum(P, W, H, 0)
    a = {k: v.to(dev).requires_grad_(True) for k, v in scenes.activated(params).items()}
    view, proj, campos = cam.matrices()
    t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)
    feats = torch.cat([a["means3D"][:, 2:3], torch.ones_like(a["means3D"][:, :1]), a["means3D"][:, 2:3] ** 2], 1).detach().requires_grad_(True)
    rs3 = splatam.Camera(H, W, cam.tanfovx, cam.tanfovy, torch.zeros(3, device=dev), 1.0, t(view).unsqueeze(0),
                         t(proj).unsqueeze(0), 0, t(campos), False, False)
    rs6 = rs3._replace(bg=torch.zeros(6, device=dev))
    dL = torch.from_numpy(np.random.default_rng(100).uniform(-1, 1, (6, H, W)).astype(np.float32)).to(dev)
    m2d = torch.zeros_like(a["means3D"])
    leaves = [a["means3D"], a["colors_precomp"], feats, a["opacities"], a["scales"], a["rotations"]]

    def ours(i):
        c6, _, _, _ = GaussianRasterizer(rs6)(means3D=a["means3D"], means2D=m2d, opacities=a["opacities"],
                                              colors_precomp=torch.cat([a["colors_precomp"], feats], 1),
                                              scales=a["scales"], rotations=a["rotations"])
        torch.autograd.grad((c6 * dL).sum(), leaves)

    def ref(i):
        R = RefRenderer(rs3)
        kw = dict(means3D=a["means3D"], means2D=m2d, opacities=a["opacities"], scales=a["scales"], rotations=a["rotations"])
        im, _, _, _ = R(colors_precomp=a["colors_precomp"], **kw)
        ds, _, _, _ = R(colors_precomp=feats, **kw)
        torch.autograd.grad((im * dL[:3]).sum() + (ds * dL[3:]).sum(), leaves)

    res = {"config": "cfg1: S-frustum 100k Gaussians, 640x480, one RGB-D+sil bundle fwd+bwd (rasterizer API, autograd)"}
    ms = timeit(ours, max(args.steps, 100), max(args.warmup, 10), dev)  # host-bound path: average out the jitter
    res["ours"] = {"ms": ms, "mpix_s": W * H / ms / 1e3}
    capi.lib().as_profile_enable(1)
    capi.profile_read()
    t0 = time.time()
    for i in range(10):
        ours(i)
    torch.cuda.synchronize(dev)
    res["ours"]["wall_ms_with_stage_events"] = (time.time() - t0) * 100
    st = capi.profile_read()
    capi.lib().as_profile_enable(0)
    res["ours"]["stage_ms"] = {k: round(v[0] / 10, 4) for k, v in st.items() if v[1]}
    if RefRenderer:
        ms = timeit(ref, max(args.steps, 100), max(args.warmup, 10), dev)
        res["ref_cuda"] = {"ms": ms, "mpix_s": W * H / ms / 1e3}
    return res


def cfg3(dev, RefRenderer, args):
    """Mapping: 60 iterations fwd+bwd+Adam on 1M Gaussians over one keyframe."""
    from oracle import splatam_oracle as so
    W, H, P, iters = 640, 480, 1_000_000, 60
    params_cpu, cams = scenes.s_room(P, 4, seed=3, W=W, H=H)
    rots, trans = scenes.camera_track(cams)
    gt = gt_frames(params_cpu, cams[:1], dev, seed=13)[0]
    cam = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device=dev)
    cur = {"cam": cam, "im": gt[:3], "depth": gt[3:4], "w2c": torch.eye(4, device=dev)}

    def fresh():
        p = {k: v.clone().to(dev).requires_grad_(True) for k, v in params_cpu.items()}
        p["cam_unnorm_rots"] = rots.clone().to(dev).requires_grad_(True)
        p["cam_trans"] = trans.clone().to(dev).requires_grad_(True)
        return p

    res = {"config": f"cfg3: S-room {P} Gaussians, {W}x{H}, {iters} mapping iterations (get_loss(mapping) + backward + Adam)"}
    # ours
    p = fresh()
    opt = FusedAdam(p, MAP_LRS)
    variables = {"max_2D_radius": torch.zeros(P, device=dev)}
    ws = workspace_for(p, cam)

    def ours(i):
        loss, _, _ = splatam.get_loss(p, cur, variables, 0, LW, False, 0.98, True, False, mapping=True)
        loss.backward()
        opt.step(ws.grad_params)
        for k in PARAM_ORDER:
            p[k].grad = None

    ms = timeit(ours, iters, 5, dev)
    n_inst, over = ws.last_status()
    assert not over
    from activesplat_b200 import capi
    capi.lib().as_profile_enable(1)
    capi.profile_read()
    for i in range(10):
        ours(i)
    st = capi.profile_read()
    capi.lib().as_profile_enable(0)
    res["ours"] = {"ms_per_iter": ms, "mpix_s": W * H / ms / 1e3, "instances": n_inst,
                   "stage_ms": {k: round(v[0] / 10, 4) for k, v in st.items() if v[1]}}
    if RefRenderer:
        p2 = fresh()
        opt2 = torch.optim.Adam([{"params": [v], "name": k, "lr": MAP_LRS[k]} for k, v in p2.items()], lr=0.0, eps=1e-15)

        def ref(i):
            loss, _, _ = so.get_loss(p2, cur, 0, LW, False, 0.98, True, False, mapping=True, Renderer=RefRenderer)
            loss.backward()
            opt2.step()
            opt2.zero_grad(set_to_none=True)

        ms = timeit(ref, iters, 5, dev)
        res["ref_cuda"] = {"ms_per_iter": ms, "mpix_s": W * H / ms / 1e3}
        # a second, identical run of the reference: its float atomics are unordered, and Adam with eps 1e-15
        # turns every noise-level gradient into a full-size step, so the reference does not reproduce itself
        p3 = fresh()
        opt3 = torch.optim.Adam([{"params": [v], "name": k, "lr": MAP_LRS[k]} for k, v in p3.items()], lr=0.0, eps=1e-15)

        def ref_again(i):
            loss, _, _ = so.get_loss(p3, cur, 0, LW, False, 0.98, True, False, mapping=True, Renderer=RefRenderer)
            loss.backward()
            opt3.step()
            opt3.zero_grad(set_to_none=True)

        timeit(ref_again, iters, 5, dev)
        res["ref_vs_ref_rel_l2"] = {k: float((p3[k].detach() - p2[k].detach()).norm() /
                                             ((p2[k].detach() - params_cpu[k].to(dev)).norm() + 1e-12)) for k in PARAM_ORDER}
        # parameter drift between the two after the same number of steps (both started from the same map)
        drift = {}
        for k in PARAM_ORDER:
            a_, b_ = p[k].detach(), p2[k].detach()
            upd = b_ - params_cpu[k].to(dev)
            drift[k] = float((a_ - b_).norm() / (upd.norm() + 1e-12))
        res["rel_l2_param_diff_over_update"] = drift
    return res


def cfg4(dev, RefRenderer, args, rank, world):
    """GetOpacity: 256 candidate views of a 1M-Gaussian map, forward only, views sharded over ranks."""
    from oracle import splatam_oracle as so
    P, n_views = 1_000_000, 256
    params_cpu, _ = scenes.s_room(P, 1, seed=4)
    params = {k: v.to(dev) for k, v in params_cpu.items()}
    res = {"config": f"cfg4: S-room {P} Gaussians, {n_views} candidate views, forward only, {world} GPU(s)"}
    for tag, (w, h, mk) in {"120x150": (120, 150, None), "640x480": (640, 480, "big")}.items():
        views = scenes.candidate_views(n_views, seed=4)
        if mk:
            fx, fy, cx, cy = scenes.compute_intrinsics(w, h, math.radians(90.0))
            views = [scenes.CameraSpec(w, h, fx, fy, cx, cy, v.w2c) for v in views]
        n_streams = min(32, max(1, -(-n_views // world)))  # every view of the shard in flight: low-res views are latency-bound

        def run_once(_i):
            return asd.score_candidate_views(params, views, n_streams=n_streams)

        run_once(0)
        torch.cuda.synchronize(dev)
        if world > 1:
            torch.distributed.barrier()
        t0 = time.time()
        scores = run_once(1)
        torch.cuda.synchronize(dev)
        if world > 1:
            torch.distributed.barrier()
        dt = time.time() - t0
        if world > 1:
            tt = torch.tensor([dt], device=dev)
            torch.distributed.all_reduce(tt, op=torch.distributed.ReduceOp.MAX)
            dt = float(tt.item())
        res[tag] = {"ours": {"views_per_s": n_views / dt, "mpix_s": n_views * w * h / dt / 1e6,
                             "mean_invisibility": float(scores[:, 0].mean()), "streams": n_streams}}
        asd._VIEW_WORKSPACES.clear()  # 32 workspaces of a 1M-Gaussian map are ~15 GB
        if RefRenderer and rank == 0 and world == 1:
            def ref_view(v):
                c = views[v]
                cam = splatam.setup_camera(c.width, c.height, c.k(), c.w2c, device=dev)
                w2c = torch.as_tensor(c.w2c).to(dev).float()
                rv, drv = so.get_rendervars(params, w2c)       # re-activates all P parameters per view
                _, _, opacity, _ = so.render(cam, splatam.with_background(cam, [1.0, 1.0, 1.0]), rv, drv, Renderer=RefRenderer)
                return torch.stack([(1.0 - opacity).sum(), opacity.mean()])

            n_ref = 32
            for v in range(4):
                ref_view(v)
            torch.cuda.synchronize(dev)
            t0 = time.time()
            rs = [ref_view(v) for v in range(n_ref)]
            torch.cuda.synchronize(dev)
            dt = time.time() - t0
            res[tag]["ref_cuda"] = {"views_per_s": n_ref / dt, "mpix_s": n_ref * w * h / dt / 1e6, "views_timed": n_ref,
                                    "max_score_diff": float((torch.stack(rs)[:, 0] - scores[:n_ref, 0]).abs().max())}
    return res


def cfg5(dev, args, rank, world):
    """MP3D-scale mapping: 3M Gaussians, 64-keyframe batch split over ranks.  The exchange + update step is timed two
    ways: NCCL all-reduce of the flat gradient followed by the fused Adam on every rank, and the single kernel that
    reduces over peer memory, updates the owned slab and writes it to every rank (dist.PeerShardedAdam)."""
    W, H, P, K = 640, 480, 3_000_000, 64
    params_cpu, cams = scenes.s_room(P, K, seed=5, W=W, H=H)
    rots, trans = scenes.camera_track(cams)
    cam = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device=dev)
    mine = asd.shard_indices(K, rank, world)
    gts = dict(zip(mine, gt_frames(params_cpu, [cams[k] for k in mine], dev, seed=15)))
    eye = torch.eye(4, device=dev)
    out = {}
    for kind in ("allreduce", "peer"):
        p = {k: v.to(dev) for k, v in params_cpu.items()}
        p["cam_unnorm_rots"], p["cam_trans"] = rots.to(dev), trans.to(dev)
        ws = workspace_for(p, cam)
        opt = FusedAdam(p, MAP_LRS) if kind == "allreduce" else asd.PeerShardedAdam(p, MAP_LRS, ws.grad_params)
        t_ex = []

        def frame_grad_fn(k, accumulate):
            with torch.no_grad():
                ws.forward(p, cam, k, sil_w2c=eye)
                ws.mapping_loss(gts[k][:3], gts[k][3:4], LW["im"], LW["depth"])
                ws.backward(gaussians_grad=True, camera_grad=False, accumulate=accumulate)
            return ws.losses[0].clone()

        def step(_i):
            e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            loss = None
            for j, k in enumerate(mine):
                l = frame_grad_fn(k, j > 0)
                loss = l if loss is None else loss + l
            e0.record()
            if kind == "allreduce":
                asd.allreduce_gradients(ws.grad_params)
                e1.record()
                opt.step(ws.grad_params)
            else:
                e1.record()
                opt.step()
            e2.record()
            t_ex.append((e0, e1, e2))
            return loss

        for i in range(2):
            step(i)
        torch.cuda.synchronize(dev)
        if world > 1:
            torch.distributed.barrier()
        t_ex.clear()
        n = 3
        t0 = time.time()
        for i in range(n):
            step(i)
        torch.cuda.synchronize(dev)
        if world > 1:
            torch.distributed.barrier()
        dt = (time.time() - t0) / n
        ex_ms = float(np.mean([a.elapsed_time(c) for a, _, c in t_ex]))
        ar_ms = float(np.mean([a.elapsed_time(b) for a, b, _ in t_ex]))
        if world > 1:
            tt = torch.tensor([dt, ex_ms], device=dev)
            torch.distributed.all_reduce(tt, op=torch.distributed.ReduceOp.MAX)
            dt, ex_ms = float(tt[0]), float(tt[1])
        R, over = ws.last_status()
        assert not over
        out[kind] = {"s_per_step": dt, "mpix_s": K * W * H / dt / 1e6, "exchange_plus_adam_ms": ex_ms,
                     "keyframes_per_rank": len(mine)}
        if kind == "allreduce":
            out[kind]["allreduce_ms"] = ar_ms
        chk = float(p["means3D"].double().sum())
        out[kind]["means3D_checksum"] = chk
        del opt, p
        from activesplat_b200 import frame as _fr
        _fr._WORKSPACES.clear()
    return {"config": f"cfg5: S-room {P} Gaussians, {K}-keyframe batch over {world} GPU(s); exchange of the [P,14] fp32 "
                      f"gradient ({P * 14 * 4 / 1e6:.0f} MB) + Adam per step, by NCCL all-reduce + fused Adam and by the "
                      "peer-memory reduce+Adam kernel",
            "ours": out["allreduce"], "allreduce_path": out["allreduce"], "peer_path": out["peer"]}


def scoring(dev):
    """The step after the renders: get_invisibility_clusters on 256 down-sampled panoramas
    (75x180), one as_dbscan_grid launch for the batch, next to the reference's CPU path (sklearn DBSCAN per image)."""
    from activesplat_b200 import scoring as sc
    from oracle import scoring_oracle as so
    n = 256
    imgs = np.stack([so.blob_image(s) for s in range(n)])
    batch = torch.from_numpy(imgs).to(dev)
    for _ in range(3):
        sc.dbscan_grid(batch, 0.3, 5.0, 10)
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        labels, ncl, si, sv = sc.dbscan_grid(batch, 0.3, 5.0, 10)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / 10
    t0 = time.time()
    n_cpu = 16
    for s in range(n_cpu):
        so.get_invisibility_clusters(imgs[s], 30)
    cpu_ms = (time.time() - t0) * 1e3 / n_cpu
    ok = all(np.array_equal(labels[s].cpu().numpy(), so.dbscan_labels_image(imgs[s], 0.3, 5, 10)) for s in range(n_cpu))
    return {"config": f"scoring: DBSCAN clusters of {n} panoramas 75x180 (eps 5, min_samples 10, invisibility > 0.3)",
            "ours": {"ms_per_batch": ms, "images_per_s": n / ms * 1e3},
            "cpu_sklearn": {"ms_per_image": cpu_ms, "images_per_s": 1e3 / cpu_ms, "images_timed": n_cpu, "cores": os.cpu_count()},
            "labels_identical_on_checked_images": bool(ok)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="1,3,4")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.distributed.init_process_group("nccl", device_id=dev)
    RefRenderer = load_ref_renderer()
    out = {"n_gpus": world}
    for c in [int(x) for x in args.configs.split(",")]:
        if c == 1 and rank == 0 and world == 1:
            out["cfg1"] = cfg1(dev, RefRenderer, args)
        elif c == 3 and rank == 0 and world == 1:
            out["cfg3"] = cfg3(dev, RefRenderer, args)
        elif c == 4:
            out["cfg4"] = cfg4(dev, RefRenderer, args, rank, world)
        elif c == 5:
            out["cfg5"] = cfg5(dev, args, rank, world)
        elif c == 6 and rank == 0 and world == 1:
            out["cfg6"] = scoring(dev)
        if rank == 0:
            print(json.dumps({k: v for k, v in out.items() if k == f"cfg{c}"}), flush=True)
    if rank == 0 and args.out:
        with open(args.out, "w") as f:
            json.dump(out, f, indent=1)
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
