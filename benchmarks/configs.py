#!/usr/bin/env python
"""The five BASELINE.json configs, ours next to the reference pipeline on the same GPU.

  python benchmarks/configs.py [--configs 1,3,4] [--out gpurun_out/configs.json]
  torchrun --nproc-per-node N benchmarks/configs.py --configs 4,5      (multi-GPU legs)

"reference pipeline" = the reference's torch code restated (oracle/splatam_oracle.py) driving the
UNMODIFIED reference CUDA rasterizer built for sm_100a (oracle/_ref): two raster passes per frame,
torch-eager glue, torch.optim.Adam -- what ActiveSplat runs today.  It is a measured baseline, not a
product path.  bench.py is the contract benchmark (cfg3 headline, the others as sub-results through the
functions below); run directly, this script gives the long form of each config (cfg3 with the 60-step parameter drift).
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
from activesplat_b200.frame import PARAM_ORDER, FusedAdam, gradients_stay_in_workspace, workspace_for  # noqa: E402

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


TRACK_LRS = dict(means3D=0.0, rgb_colors=0.0, unnorm_rotations=0.0, logit_opacities=0.0, log_scales=0.0,
                 cam_unnorm_rots=0.001, cam_trans=0.004)  # config/splatam/online_habitat_sim.py:37-45


def load_ref_python(contiguous_quat=False):
    """The reference's own SplaTAM-layer Python (byte-compiled into oracle/_ref/refpy by oracle/build_ref_py.py) on
    the reference's own CUDA rasterizer; None when either is absent.  contiguous_quat=True wraps (does not edit) its
    quat_mult so that the rasterizer receives row-major rotations (see load_ref_renderer)."""
    from oracle import build_ref_py
    if not (build_ref_py.available() and os.path.exists(os.path.join(ROOT, "oracle", "_ref", "ref_dgr_C.so"))):
        return None
    ns = build_ref_py.load(build_ref_py.load_ref_extension())
    if contiguous_quat:
        orig = ns.slam_helpers.quat_mult
        ns.slam_helpers.quat_mult = lambda a, b: orig(a, b).contiguous()
    return ns


def ref_frame(ns, W, H, k, gt, dev, t_idx=0):
    """curr_data / variables as the reference's mapper builds them (src/mapper/splatam/__init__.py:354-397)."""
    cam = ns.recon_helpers.setup_camera(W, H, k, np.eye(4))
    curr = {"cam": cam, "im": gt[:3], "depth": gt[3:4], "id": t_idx, "w2c": torch.eye(4, device=dev),
            "intrinsics": torch.tensor(k).float().to(dev)}
    return curr


def ref_variables(P, dev):
    return {"max_2D_radius": torch.zeros(P, device=dev), "means2D_gradient_accum": torch.zeros(P, device=dev),
            "denom": torch.zeros(P, device=dev), "timestep": torch.zeros(P, device=dev)}


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
    """RGB-D 'ground truth' frames: a perturbed copy of the map (seed None: `params_cpu` is already that copy)
    rendered at the given poses."""
    src = params_cpu if seed is None else scenes.perturbed(params_cpu, seed=seed)
    gtp = {k: v.to(dev) for k, v in src.items()}
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


def _pose_error(rots_a, trans_a, rots_b, trans_b, t):
    """(rotation angle in degrees, translation distance in mm) between two poses stored SplaTAM-style."""
    qa = torch.nn.functional.normalize(rots_a[0, :, t].detach().double().cpu(), dim=0)
    qb = torch.nn.functional.normalize(rots_b[0, :, t].detach().double().cpu(), dim=0)
    ang = 2 * torch.acos(torch.clamp(torch.abs((qa * qb).sum()), max=1.0))
    dt = trans_a[0, :, t].detach().double().cpu() - trans_b[0, :, t].detach().double().cpu()
    return float(torch.rad2deg(ang)), float(dt.norm() * 1e3)


def cfg2(dev, with_ref, args):
    """Camera tracking: 40 pose-only iterations (get_loss(tracking=True) + backward + Adam on cam_unnorm_rots / cam_trans,
    lr 1e-3 / 4e-3) on a 500k-Gaussian map at 640x480, pose initialised 2 cm / 1 degree off; final pose vs the
    reference's own loop (its get_loss + initialize_optimizer(tracking=True)) from the same start."""
    from activesplat_b200.frame import FusedPoseAdam, GraphedStep
    W, H, P, iters, t = 640, 480, 500_000, 40, 1
    params_cpu, cams = scenes.s_room(P, 4, seed=1, W=W, H=H)
    rots_gt, trans_gt = scenes.camera_track(cams)
    gt = gt_frames(params_cpu, [cams[t]], dev, seed=11)[0]
    # initial pose: ground truth perturbed by 2 cm and 1 degree
    g = torch.Generator().manual_seed(2)
    axis = torch.nn.functional.normalize(torch.randn(3, generator=g), dim=0)
    half = math.radians(1.0) / 2
    dq = torch.cat([torch.tensor([math.cos(half)]), math.sin(half) * axis])
    q = rots_gt[0, :, t]
    q0 = torch.stack([dq[0] * q[0] - dq[1] * q[1] - dq[2] * q[2] - dq[3] * q[3],
                      dq[0] * q[1] + dq[1] * q[0] + dq[2] * q[3] - dq[3] * q[2],
                      dq[0] * q[2] - dq[1] * q[3] + dq[2] * q[0] + dq[3] * q[1],
                      dq[0] * q[3] + dq[1] * q[2] - dq[2] * q[1] + dq[3] * q[0]])
    rots0, trans0 = rots_gt.clone(), trans_gt.clone()
    rots0[0, :, t] = q0
    trans0[0, :, t] += 0.02 * torch.nn.functional.normalize(torch.randn(3, generator=g), dim=0)
    k = cams[0].k()

    def fresh():
        p = {k_: v.clone().to(dev).requires_grad_(True) for k_, v in params_cpu.items()}
        p["cam_unnorm_rots"] = rots0.clone().to(dev).requires_grad_(True)
        p["cam_trans"] = trans0.clone().to(dev).requires_grad_(True)
        return p

    res = {"config": f"cfg2: S-room {P} Gaussians, {W}x{H}, {iters} pose-only iterations (get_loss(tracking) + backward + "
                     "Adam on the camera pose), start 2 cm / 1 deg off"}
    start = _pose_error(rots0, trans0, rots_gt, trans_gt, t)
    res["start_error_deg_mm"] = start
    # ---- ours ----
    cam = splatam.setup_camera(W, H, k, np.eye(4), device=dev)
    cur = {"cam": cam, "im": gt[:3], "depth": gt[3:4], "w2c": torch.eye(4, device=dev)}

    def run_ours(graph):
        p = fresh()
        opt = FusedPoseAdam(p, TRACK_LRS)
        variables = {"max_2D_radius": torch.zeros(P, device=dev)}

        def it():
            for v in p.values():
                v.grad = None
            loss, _, _ = splatam.get_loss(p, cur, variables, t, LW, True, 0.99, True, False, tracking=True)
            loss.backward()
            opt.step()
            return loss

        it()  # sizes the workspace (one synchronising look at the instance count); then start over
        p2 = fresh()
        with torch.no_grad():
            for k_ in p:
                p[k_].copy_(p2[k_])
        opt.exp_avg.zero_(), opt.exp_avg_sq.zero_(), opt.step_state.zero_()
        step = it
        if graph:
            gs = GraphedStep(it, warmup=0)     # capture runs nothing; the first replay is iteration 1
            step = gs
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            loss = step()
        e1.record()
        torch.cuda.synchronize(dev)
        return p, e0.elapsed_time(e1) / iters, float(loss.detach())

    p_e, ms_e, loss_e = run_ours(False)
    res["ours"] = {"ms_per_iter_eager": ms_e, "final_loss": loss_e,
                   "final_error_vs_gt_deg_mm": _pose_error(p_e["cam_unnorm_rots"], p_e["cam_trans"], rots_gt, trans_gt, t)}
    try:
        p_g, ms_g, loss_g = run_ours(True)
        res["ours"].update(ms_per_iter=ms_g, mpix_s=W * H / ms_g / 1e3, final_loss_graph=loss_g,
                           graph_vs_eager_deg_mm=_pose_error(p_g["cam_unnorm_rots"], p_g["cam_trans"],
                                                             p_e["cam_unnorm_rots"], p_e["cam_trans"], t))
    except Exception as ex:
        res["ours"]["graph_failed"] = f"{type(ex).__name__}: {ex}"[:200]
        res["ours"].update(ms_per_iter=ms_e, mpix_s=W * H / ms_e / 1e3)
    # ---- the reference's own loop ----
    ns = load_ref_python() if with_ref else None
    if ns is not None:
        def run_ref():
            p = fresh()
            p = {k_: torch.nn.Parameter(v.detach()) for k_, v in p.items()}
            opt = ns.splatam.initialize_optimizer(p, TRACK_LRS, tracking=True)
            curr = ref_frame(ns, W, H, k, gt, dev, t)
            variables = ref_variables(P, dev)
            torch.cuda.synchronize(dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                loss, variables, _ = ns.splatam.get_loss(p, curr, variables, t, LW, True, 0.99, True, False, tracking=True)
                loss.backward()
                with torch.no_grad():
                    opt.step()
                    opt.zero_grad(set_to_none=True)
            e1.record()
            torch.cuda.synchronize(dev)
            return p, e0.elapsed_time(e1) / iters, float(loss)

        run_ref()  # warm-up (allocator, cuDNN-free but lazy module loads)
        p_r, ms_r, loss_r = run_ref()
        p_r2, _, _ = run_ref()
        res["ref_cuda"] = {"ms_per_iter": ms_r, "mpix_s": W * H / ms_r / 1e3, "final_loss": loss_r,
                           "final_error_vs_gt_deg_mm": _pose_error(p_r["cam_unnorm_rots"], p_r["cam_trans"], rots_gt, trans_gt, t)}
        res["final_pose_ours_vs_reference_deg_mm"] = _pose_error(p_e["cam_unnorm_rots"], p_e["cam_trans"],
                                                                 p_r["cam_unnorm_rots"], p_r["cam_trans"], t)
        res["final_pose_reference_vs_its_second_run_deg_mm"] = _pose_error(p_r2["cam_unnorm_rots"], p_r2["cam_trans"],
                                                                           p_r["cam_unnorm_rots"], p_r["cam_trans"], t)
    return res


def cfg3(dev, RefRenderer, args, iters=60):
    """Mapping: 60 iterations fwd+bwd+Adam on 1M Gaussians over one keyframe; time per iteration and the parameter
    drift after EXACTLY `iters` steps on both sides, against the reference's own loop (stock, and fed row-major
    rotations) and against a second run of the reference itself."""
    W, H, P = 640, 480, 1_000_000
    params_cpu, cams = scenes.s_room(P, 4, seed=3, W=W, H=H)
    rots, trans = scenes.camera_track(cams)
    gt = gt_frames(params_cpu, cams[:1], dev, seed=13)[0]
    k = cams[0].k()
    cam = splatam.setup_camera(W, H, k, np.eye(4), device=dev)
    cur = {"cam": cam, "im": gt[:3], "depth": gt[3:4], "w2c": torch.eye(4, device=dev)}

    def fresh():
        p = {k_: v.clone().to(dev).requires_grad_(True) for k_, v in params_cpu.items()}
        p["cam_unnorm_rots"] = rots.clone().to(dev).requires_grad_(True)
        p["cam_trans"] = trans.clone().to(dev).requires_grad_(True)
        return p

    res = {"config": f"cfg3: S-room {P} Gaussians, {W}x{H}, {iters} mapping iterations (get_loss(mapping) + backward + Adam)"}

    def run_ours(n, profile=False):
        p = fresh()
        opt = FusedAdam(p, MAP_LRS)
        variables = {"max_2D_radius": torch.zeros(P, device=dev)}
        ws = workspace_for(p, cam)

        def it(i):
            for k_ in PARAM_ORDER:
                p[k_].grad = None
            loss, _, _ = splatam.get_loss(p, cur, variables, 0, LW, False, 0.98, True, False, mapping=True)
            with gradients_stay_in_workspace():
                loss.backward()
            opt.step(ws.grad_params)

        if profile:
            capi.lib().as_profile_enable(1)
            capi.profile_read()
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(n):
            it(i)
        e1.record()
        torch.cuda.synchronize(dev)
        st = None
        if profile:
            st = capi.profile_read()
            capi.lib().as_profile_enable(0)
        return p, e0.elapsed_time(e1) / n, st, ws.last_status()

    run_ours(5)                                   # warm-up on a throw-away copy
    p, ms, _, (n_inst, over) = run_ours(iters)    # exactly `iters` steps from the initial map
    assert not over
    _, _, st, _ = run_ours(10, profile=True)      # stage attribution, again on its own copy
    res["ours"] = {"ms_per_iter": ms, "mpix_s": W * H / ms / 1e3, "instances": n_inst,
                   "stage_ms": {k_: round(v[0] / 10, 4) for k_, v in st.items() if v[1]}}

    def run_ref(ns, n):
        p2 = {k_: torch.nn.Parameter(v.detach()) for k_, v in fresh().items()}
        opt2 = ns.splatam.initialize_optimizer(p2, MAP_LRS, tracking=False)
        curr = ref_frame(ns, W, H, k, gt, dev, 0)
        variables = ref_variables(P, dev)
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            loss, variables, _ = ns.splatam.get_loss(p2, curr, variables, 0, LW, False, 0.98, True, False, mapping=True)
            loss.backward()
            with torch.no_grad():
                opt2.step()
                opt2.zero_grad(set_to_none=True)
        e1.record()
        torch.cuda.synchronize(dev)
        return p2, e0.elapsed_time(e1) / n

    def rel_l2(a, b):
        """|a - b| / |b - initial| per tensor: the difference between two runs in units of the update itself."""
        return {k_: float((a[k_].detach() - b[k_].detach()).norm() /
                          ((b[k_].detach() - params_cpu[k_].to(dev)).norm() + 1e-12)) for k_ in PARAM_ORDER}

    ns_stock = load_ref_python() if RefRenderer is not None else None
    if ns_stock is not None:
        run_ref(ns_stock, 5)
        p_stock, ms_ref = run_ref(ns_stock, iters)
        p_stock2, _ = run_ref(ns_stock, iters)
        ns_c = load_ref_python(contiguous_quat=True)
        p_c, _ = run_ref(ns_c, iters)
        p_c2, _ = run_ref(ns_c, iters)
        res["ref_cuda"] = {"ms_per_iter": ms_ref, "mpix_s": W * H / ms_ref / 1e3}
        res["drift_after_%d_steps" % iters] = {
            "unit": "|a - b| / |b - initial map|, per tensor (L2)",
            "ours_vs_reference_stock": rel_l2(p, p_stock),
            "ours_vs_reference_fed_row_major_rotations": rel_l2(p, p_c),
            "reference_vs_its_second_run_stock": rel_l2(p_stock2, p_stock),
            "reference_vs_its_second_run_row_major": rel_l2(p_c2, p_c),
            "note": "the stock reference backward misreads SplaTAM's column-major rotations (DESIGN.md 2.1), so scales and "
                    "rotations drift from it by construction; the row-major variant is the like-for-like comparison"}
    return res


def cfg4(dev, RefRenderer, args, rank, world, params=None):
    """GetOpacity: 256 candidate views of a 1M-Gaussian map, forward only, views sharded over ranks."""
    from oracle import splatam_oracle as so
    P, n_views = 1_000_000, 256
    if params is None:
        params_cpu, _ = scenes.s_room(P, 1, seed=4)
        params = {k: v.to(dev) for k, v in params_cpu.items()}
    P = int(params["means3D"].shape[0])
    res = {"config": f"cfg4: S-room {P} Gaussians, {n_views} candidate views, forward only, {world} GPU(s)"}
    for tag, (w, h, mk) in {"120x150": (120, 150, None), "640x480": (640, 480, "big")}.items():
        views = scenes.candidate_views(n_views, seed=4)
        if mk:
            fx, fy, cx, cy = scenes.compute_intrinsics(w, h, math.radians(90.0))
            views = [scenes.CameraSpec(w, h, fx, fy, cx, cy, v.w2c) for v in views]
        n_streams = min(32, max(1, -(-n_views // world)))  # every view of the shard in flight: low-res views are latency-bound

        def run_once(_i):
            return asd.score_candidate_views(params, views, n_streams=n_streams)

        # three sweeps to warm up: with one view per workspace and sweep (8 GPUs x 32 streams) a workspace needs two frames
        # before it has seen an instance count, switches to the chunked blend and allocates that path's scratch
        for i in range(3):
            run_once(i)
        torch.cuda.synchronize(dev)
        if world > 1:
            torch.distributed.barrier()
        n_sweeps = 3
        t0 = time.time()
        for i in range(n_sweeps):
            scores = run_once(1 + i)
        torch.cuda.synchronize(dev)
        if world > 1:
            torch.distributed.barrier()
        dt = (time.time() - t0) / n_sweeps
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
    # A whole GLOBAL GetOpacity query (visualizer.py:2183-2195 + :991-1001): 86 Voronoi nodes x 3 look-around views,
    # panoramas, DBSCAN clusters and convex-hull volumes per node -- renders AND scoring, nodes sharded over the ranks.
    from activesplat_b200 import scoring
    nodes = scenes.candidate_views(86 * 3, seed=4)[::3]
    positions = np.stack([np.linalg.inv(np.asarray(v.w2c))[:3, 3] for v in nodes])
    c2w = np.eye(4)
    scoring.get_global_invisibility(params, c2w, positions)
    torch.cuda.synchronize(dev)
    if world > 1:
        torch.distributed.barrier()
    t0 = time.time()
    for _ in range(3):
        inv, vol = scoring.get_global_invisibility(params, c2w, positions)
    torch.cuda.synchronize(dev)
    if world > 1:
        torch.distributed.barrier()
    dt = (time.time() - t0) / 3
    if world > 1:
        tt = torch.tensor([dt], device=dev)
        torch.distributed.all_reduce(tt, op=torch.distributed.ReduceOp.MAX)
        dt = float(tt.item())
    res["global_query_86_nodes"] = {"ms_per_query": dt * 1e3, "nodes_with_volume": int((vol > 0).sum()),
                                    "sum_volume": float(vol.sum()),
                                    "what": "scoring.get_global_invisibility: 258 renders + per-node DBSCAN + hull volume"}
    asd._VIEW_WORKSPACES.clear()
    return res


def cfg5(dev, args, rank, world, with_peer=True):
    """MP3D-scale mapping: 3M Gaussians, 64-keyframe batch split over ranks (dist.mapping_step_keyframe_batch).  After a
    first round-robin step the keyframes are dealt by their measured Gaussian-tile instance counts, so that no rank
    waits for the one that drew the heavy views.  The exchange + update is timed two ways: NCCL all-reduce of the flat
    gradient followed by the fused Adam on every rank, and the single kernel that reduces over peer memory, updates the
    owned slab and writes it to every rank (dist.PeerShardedAdam)."""
    W, H, P, K = 640, 480, 3_000_000, 64
    params_cpu, cams = scenes.s_room(P, K, seed=5, W=W, H=H)
    rots, trans = scenes.camera_track(cams)
    cam = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device=dev)
    gts = gt_frames(params_cpu, cams, dev, seed=15)      # every rank holds every keyframe's RGB-D frame (315 MB)
    eye = torch.eye(4, device=dev)
    out = {}
    for kind in (("allreduce", "peer") if with_peer else ("allreduce",)):
        p = {k: v.to(dev) for k, v in params_cpu.items()}
        p["cam_unnorm_rots"], p["cam_trans"] = rots.to(dev), trans.to(dev)
        ws = workspace_for(p, cam)
        opt = FusedAdam(p, MAP_LRS) if kind == "allreduce" else asd.PeerShardedAdam(p, MAP_LRS, ws.grad_params)
        inst = torch.zeros(K, dtype=torch.int32, device=dev)
        ev = {}

        def frame_grad_fn(k, accumulate):
            with torch.no_grad():
                ws.forward(p, cam, k, sil_w2c=eye)
                ws.mapping_loss(gts[k][:3], gts[k][3:4], LW["im"], LW["depth"])
                ws.backward(gaussians_grad=True, camera_grad=False, accumulate=accumulate)
            inst[k] = ws.status[0]
            ev["frames_done"] = torch.cuda.Event(enable_timing=True)
            ev["frames_done"].record()
            return ws.losses[0].clone()

        def optimizer_step(flat):
            ev["reduced"] = torch.cuda.Event(enable_timing=True)
            ev["reduced"].record()
            opt.step(flat) if kind == "allreduce" else opt.step()
            ev["updated"] = torch.cuda.Event(enable_timing=True)
            ev["updated"].record()

        def step(weights):
            inst.zero_()
            if kind == "allreduce":
                return asd.mapping_step_keyframe_batch(frame_grad_fn, K, ws.grad_params, optimizer_step, weights=weights)
            # the peer kernel does the exchange itself: same deal, no all-reduce
            mine = asd.shard_indices(K, rank, world) if weights is None else asd.deal_by_weight(weights, world)[rank]
            loss = None
            for j, k in enumerate(mine):
                l = frame_grad_fn(k, j > 0)
                loss = l if loss is None else loss + l
            optimizer_step(ws.grad_params)
            return loss

        def measured_weights():
            if world > 1:
                torch.distributed.all_reduce(inst)
            return inst.cpu().tolist()

        step(None)                       # round-robin: sizes the workspace and measures every keyframe's weight
        weights = measured_weights()
        step(weights)
        res = {}
        for tag, wts in (("round_robin", None), ("by_weight", weights)):
            torch.cuda.synchronize(dev)
            if world > 1:
                torch.distributed.barrier()
            n, t_ex, t_ar = 3, [], []
            t0 = time.time()
            for i in range(n):
                step(wts)
                t_ex.append((ev["frames_done"], ev["reduced"], ev["updated"]))
            torch.cuda.synchronize(dev)
            if world > 1:
                torch.distributed.barrier()
            dt = (time.time() - t0) / n
            ex_ms = float(np.mean([a.elapsed_time(c) for a, _, c in t_ex]))
            ar_ms = float(np.mean([a.elapsed_time(b) for a, b, _ in t_ex]))
            if world > 1:
                tt = torch.tensor([dt, ex_ms, ar_ms], device=dev)
                torch.distributed.all_reduce(tt, op=torch.distributed.ReduceOp.MAX)
                dt, ex_ms, ar_ms = float(tt[0]), float(tt[1]), float(tt[2])
            mine = asd.shard_indices(K, rank, world) if wts is None else asd.deal_by_weight(wts, world)[rank]
            res[tag] = {"s_per_step": dt, "mpix_s": K * W * H / dt / 1e6, "exchange_plus_adam_ms": ex_ms,
                        "keyframes_this_rank": len(mine)}
            if kind == "allreduce":
                res[tag]["allreduce_ms_incl_wait_for_slowest_rank"] = ar_ms
        R, over = ws.last_status()
        assert not over
        best = min(res.values(), key=lambda r: r["s_per_step"])
        out[kind] = dict(best, deals=res, means3D_checksum=float(p["means3D"].double().sum()),
                         instances_per_keyframe_min_max=[int(min(weights)), int(max(weights))])
        del opt, p
        from activesplat_b200 import frame as _fr
        _fr._WORKSPACES.clear()
        torch.cuda.empty_cache()
    return {"config": f"cfg5: S-room {P} Gaussians, {K}-keyframe batch over {world} GPU(s) through "
                      f"dist.mapping_step_keyframe_batch; exchange of the flat fp32 gradient, 14 floats per Gaussian ({P * 14 * 4 / 1e6:.0f} MB) "
                      "+ Adam per step inside the timed region, by NCCL all-reduce + fused Adam"
                      + (" and by the peer-memory reduce+Adam kernel" if with_peer else ""),
            "ours": out["allreduce"], **({"peer_path": out["peer"]} if with_peer else {})}


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
        elif c == 2 and rank == 0 and world == 1:
            out["cfg2"] = cfg2(dev, RefRenderer is not None, args)
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
