"""GPU: the sm_100a kernels, called through the C ABI, against the CPU oracle and the golden
vectors of the reference CUDA build.  Bit-exact for every integer / index structure; 1e-4 relative
(BASELINE.json north_star) for images and gradients."""
import ctypes
import os

import numpy as np
import pytest
import torch

import helpers

pytestmark = pytest.mark.gpu

INT_KEYS = ["radii", "tiles_touched", "keys_unsorted", "vals_unsorted", "keys", "point_list", "ranges"]
GRADS = ("means2D", "colors", "opacity", "means3D", "cov3D", "scales", "rotations")


def _check_against(out, grads, ref_out, ref_grads, n_contrib_slack):
    assert out["num_rendered"] == int(ref_out["num_rendered"])
    for k in INT_KEYS:
        helpers.compare_exact(k, out[k], ref_out[k])
    helpers.compare_exact("n_contrib", out["n_contrib"], ref_out["n_contrib"], max_mismatch=n_contrib_slack)
    vis = np.asarray(ref_out["radii"]) > 0
    helpers.compare_exact("means2D", out["means2D"][vis], np.asarray(ref_out["means2D"])[vis])
    helpers.compare_exact("depths", out["depths"][vis], np.asarray(ref_out["depths"])[vis])
    helpers.compare_float("conic_opacity", out["conic_opacity"][vis], np.asarray(ref_out["conic_opacity"])[vis])
    for k in ("color", "depth", "opacity", "final_T"):
        helpers.compare_float(k, out[k], ref_out[k])
    for k in GRADS:
        helpers.compare_grad("grad_" + k, grads[k], ref_grads[k])


@pytest.mark.parametrize("name", helpers.CASES)
def test_cuda_matches_oracle(name):
    case = helpers.make_case(name)
    st, g = helpers.run_oracle(case)
    out, grads = helpers.run_cuda(case)
    _check_against(out, grads, st, g, n_contrib_slack=max(1, int(2e-4 * st["n_contrib"].size)))
    # conic is computed with the same explicit roundings on both sides: must be identical
    vis = st["radii"] > 0
    helpers.compare_exact("conic_opacity bits", out["conic_opacity"][vis], st["conic_opacity"][vis])


@pytest.mark.parametrize("name", helpers.CASES)
def test_cuda_matches_reference_golden(name):
    path = os.path.join(helpers.GOLDEN_DIR, f"ref_{name}.npz")
    ref = np.load(path)
    case = helpers.make_case(name)
    out, grads = helpers.run_cuda(case)
    # same expf as the reference build => identical alpha decisions => n_contrib identical
    _check_against(out, grads, ref, {k: ref["grad_" + k] for k in GRADS}, n_contrib_slack=0)


@pytest.mark.parametrize("name", ["frustum_small", "lookaround_depthsil"])
def test_six_channel_bundle_equals_two_passes(name):
    """C = 6 ([rgb | z,1,z^2] in one sweep) must equal the reference's two 3-channel passes
    (src/mapper/splatam/splatam.py:208,212): same images, and gradients that are the sum of both passes."""
    case = helpers.make_case(name)
    cam_feats = helpers.depth_sil_features(torch.from_numpy(case["means3D"]), np.eye(4)).numpy()
    rgb = np.random.default_rng(5).uniform(0, 1, case["colors"].shape).astype(np.float32)
    bg6 = np.array([0.1, 0.2, 0.3, 0.0, 0.0, 0.0], np.float32)
    dL6 = np.random.default_rng(6).uniform(-1, 1, (6, case["H"], case["W"])).astype(np.float32)
    out6, g6 = helpers.run_cuda(case, colors=np.concatenate([rgb, cam_feats], 1), bg=bg6, dL=dL6)
    outa, ga = helpers.run_cuda(case, colors=rgb, bg=bg6[:3], dL=dL6[:3])
    outb, gb = helpers.run_cuda(case, colors=cam_feats, bg=bg6[3:], dL=dL6[3:])
    for k in INT_KEYS + ["n_contrib"]:
        helpers.compare_exact(k, out6[k], outa[k])
    helpers.compare_exact("color[:3]", out6["color"][:3], outa["color"])
    helpers.compare_exact("color[3:]", out6["color"][3:], outb["color"])
    helpers.compare_exact("depth", out6["depth"], outa["depth"])
    helpers.compare_exact("opacity", out6["opacity"], outa["opacity"])
    helpers.compare_grad("grad_colors[:3]", g6["colors"][:, :3], ga["colors"])
    helpers.compare_grad("grad_colors[3:]", g6["colors"][:, 3:], gb["colors"])
    for k in ("means2D", "opacity", "means3D", "scales", "rotations"):
        helpers.compare_grad("grad_" + k, g6[k], ga[k] + gb[k])
    # and against the oracle's 6-channel sweep
    st, g = helpers.run_oracle(case, colors=np.concatenate([rgb, cam_feats], 1), bg=bg6, dL=dL6)
    helpers.compare_float("color6 vs oracle", out6["color"], st["color"])
    for k in GRADS:
        helpers.compare_grad("grad6_" + k, g6[k], g[k])


def test_python_surface_autograd_matches_oracle():
    """GaussianRasterizer through torch autograd (the call the mapper makes, splatam.py:208)."""
    from activesplat_b200.rasterizer import GaussianRasterizationSettings, GaussianRasterizer
    from oracle import raster_oracle as ro
    case = helpers.make_case("frustum_small")
    outs = {}
    for dev, R in (("cuda:0", GaussianRasterizer), ("cpu", ro.GaussianRasterizer)):
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        rs = GaussianRasterizationSettings(case["H"], case["W"], case["tanfovx"], case["tanfovy"], t(case["bg"]), 1.0,
                                           t(case["view"]).unsqueeze(0), t(case["proj"]).unsqueeze(0), 0,
                                           t(case["campos"]), False, False)
        leaves = {k: t(case[k]).requires_grad_(True) for k in ("means3D", "colors", "opacities", "scales", "rotations")}
        means2D = torch.zeros_like(leaves["means3D"], requires_grad=True)
        color, radii, depth, opacity = R(rs)(means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opacities"],
                                             colors_precomp=leaves["colors"], scales=leaves["scales"],
                                             rotations=leaves["rotations"])
        assert radii.dtype == torch.int32 and tuple(color.shape) == (3, case["H"], case["W"])
        assert tuple(depth.shape) == (1, case["H"], case["W"]) and tuple(opacity.shape) == (1, case["H"], case["W"])
        (color * t(case["dL_dpix"])).sum().backward()
        outs[dev] = dict(color=color.detach().cpu().numpy(), radii=radii.cpu().numpy(),
                         means2D=means2D.grad.cpu().numpy(),
                         **{"g_" + k: v.grad.cpu().numpy() for k, v in leaves.items()})
    helpers.compare_exact("radii", outs["cuda:0"]["radii"], outs["cpu"]["radii"])
    for k in outs["cpu"]:
        if k != "radii":
            helpers.compare_float(k, outs["cuda:0"][k], outs["cpu"][k])


def test_mark_visible_and_edge_cases():
    from activesplat_b200 import _C
    from oracle import raster_oracle as ro
    case = helpers.make_case("room_partial_tiles")
    dev = "cuda:0"
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    got = _C.mark_visible(t(case["means3D"]), t(case["view"]), t(case["proj"])).cpu().numpy()
    helpers.compare_exact("mark_visible", got, ro.mark_visible(case["means3D"], case["view"]))
    e = torch.Tensor([])
    # P == 0: nothing rendered, zero outputs, empty scratch (rasterize_points.cu:83)
    R, color, radii, geom, binning, img, depth, opacity = _C.rasterize_gaussians(
        t(case["bg"]), torch.zeros((0, 3), device=dev), torch.zeros((0, 3), device=dev), torch.zeros((0, 1), device=dev),
        torch.zeros((0, 3), device=dev), torch.zeros((0, 4), device=dev), 1.0, e, t(case["view"]), t(case["proj"]),
        case["tanfovx"], case["tanfovy"], 32, 48, e, 0, t(case["campos"]), False, False)
    assert R == 0 and radii.numel() == 0 and float(color.abs().max()) == 0.0
    # every Gaussian behind the camera: R == 0, image = background, depth 15, opacity 0
    m = case["means3D"].copy()
    m[:, :] = [0, 0, -5]
    c2 = dict(case, means3D=m, view=np.eye(4, dtype=np.float32))
    out, grads = helpers.run_cuda(c2)
    assert out["num_rendered"] == 0 and (out["radii"] == 0).all()
    np.testing.assert_array_equal(out["color"], np.broadcast_to(case["bg"][:, None, None], out["color"].shape))
    assert (out["depth"] == 15.0).all() and (out["opacity"] == 0).all()
    for k in GRADS:
        assert not grads[k].any()
    # missing colours -> the reference's error message (rasterizer_impl.cu:244-247)
    with pytest.raises(RuntimeError, match="For non-RGB, provide precomputed Gaussian colors!"):
        _C.rasterize_gaussians(t(case["bg"]), t(case["means3D"]), e, t(case["opacities"]), t(case["scales"]),
                               t(case["rotations"]), 1.0, e, t(case["view"]), t(case["proj"]), case["tanfovx"],
                               case["tanfovy"], 32, 48, e, 0, t(case["campos"]), False, False)
    with pytest.raises(RuntimeError, match="means3D must have dimensions"):
        _C.rasterize_gaussians(t(case["bg"]), torch.zeros((4, 2), device=dev), e, e, e, e, 1.0, e, t(case["view"]),
                               t(case["proj"]), 1.0, 1.0, 32, 48, e, 0, t(case["campos"]), False, False)


@pytest.mark.parametrize("n,end_bit", [(1, 43), (255, 39), (2048, 43), (2049, 43), (100_000, 43), (1_000_003, 64), (50_000, 8)])
def test_sort_pairs_and_scan(n, end_bit):
    """as_sort_pairs vs a stable numpy sort (== cub::DeviceRadixSort::SortPairs semantics), and
    as_inclusive_scan_u32 vs numpy cumsum, through the C ABI with raw device pointers."""
    from activesplat_b200 import capi
    L = capi.lib()
    rng = np.random.default_rng(n)
    tiles = rng.integers(0, 1 << max(end_bit - 32, 1), n, dtype=np.uint64) if end_bit > 32 else np.zeros(n, np.uint64)
    depth = rng.uniform(0.2, 10.0, n).astype(np.float32).view(np.uint32).astype(np.uint64)
    if n > 1000:  # many exact ties to exercise stability
        depth[rng.integers(0, n, n // 2)] = depth[0]
    keys = (tiles << np.uint64(32)) | depth
    if end_bit < 64:
        keys &= np.uint64((1 << end_bit) - 1)
    vals = np.arange(n, dtype=np.uint32)
    dev = "cuda:0"
    k_in = torch.from_numpy(keys.view(np.int64)).to(dev)
    v_in = torch.from_numpy(vals.view(np.int32)).to(dev)
    k_out, v_out = torch.empty_like(k_in), torch.empty_like(v_in)
    tb = L.as_sort_temp_bytes(n)
    temp = torch.empty(tb, dtype=torch.uint8, device=dev)
    s = torch.cuda.current_stream().cuda_stream
    capi.check(L.as_sort_pairs(k_in.data_ptr(), v_in.data_ptr(), k_out.data_ptr(), v_out.data_ptr(), n, end_bit,
                               temp.data_ptr(), tb, s))
    torch.cuda.synchronize()
    order = np.argsort(keys, kind="stable")
    helpers.compare_exact("sorted keys", k_out.cpu().numpy().view(np.uint64), keys[order])
    helpers.compare_exact("sorted values", v_out.cpu().numpy().view(np.uint32), vals[order])
    helpers.compare_exact("keys_in preserved", k_in.cpu().numpy().view(np.uint64), keys)
    # scan
    x = rng.integers(0, 50, n, dtype=np.uint32)
    x_d = torch.from_numpy(x.view(np.int32)).to(dev)
    y_d = torch.empty_like(x_d)
    tot = torch.zeros(1, dtype=torch.int32, device=dev)
    sb = L.as_scan_temp_bytes(n)
    stemp = torch.empty(sb, dtype=torch.uint8, device=dev)
    capi.check(L.as_inclusive_scan_u32(x_d.data_ptr(), y_d.data_ptr(), n, tot.data_ptr(), stemp.data_ptr(), sb, s))
    torch.cuda.synchronize()
    ref = np.cumsum(x.astype(np.uint64)).astype(np.uint32)
    helpers.compare_exact("scan", y_d.cpu().numpy().view(np.uint32), ref)
    assert int(tot.item()) == int(ref[-1])


def test_full_size_properties_cfg1():
    """BASELINE config 1 size (100k Gaussians, 640x480): size-independent properties.
    sortedness + stability of the instance list, ranges partition it, offsets are the scan of tiles
    touched, per-pixel T in [0,1], opacity = 1 - T, 0 <= n_contrib <= tile list length."""
    from activesplat_b200 import scenes
    params, cam = scenes.s_frustum(100_000, 640, 480, 0)
    a = scenes.activated(params)
    view, proj, campos = cam.matrices()
    case = dict(means3D=a["means3D"].numpy(), colors=a["colors_precomp"].numpy(), opacities=a["opacities"].numpy(),
                scales=a["scales"].numpy(), rotations=a["rotations"].numpy(), scale_modifier=1.0, view=view, proj=proj,
                campos=campos, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, H=480, W=640, bg=np.zeros(3, np.float32),
                dL_dpix=np.random.default_rng(100).uniform(-1, 1, (3, 480, 640)).astype(np.float32))
    out, grads = helpers.run_cuda(case)
    k = out["keys"]
    assert (k[1:] >= k[:-1]).all()
    ties = k[1:] == k[:-1]
    assert (out["point_list"][1:][ties] > out["point_list"][:-1][ties]).all()
    helpers.compare_exact("offsets == cumsum(tiles_touched)", out["offsets"],
                          np.cumsum(out["tiles_touched"].astype(np.uint64)).astype(np.uint32))
    assert int(out["offsets"][-1]) == out["num_rendered"]
    assert ((out["radii"] > 0) == (out["tiles_touched"] > 0)).all()
    r = out["ranges"].astype(np.int64)
    assert int((r[:, 1] - r[:, 0]).sum()) == out["num_rendered"]
    tile_of_key = (k >> np.uint64(32)).astype(np.int64)
    nz = np.where(r[:, 1] > r[:, 0])[0]
    assert (tile_of_key[r[nz, 0]] == nz).all() and (tile_of_key[r[nz, 1] - 1] == nz).all()
    assert (out["final_T"] >= 0).all() and (out["final_T"] <= 1).all()
    np.testing.assert_array_equal(out["opacity"][0], np.float32(1) - out["final_T"])
    tile_len = (r[:, 1] - r[:, 0]).reshape(30, 40)
    assert (out["n_contrib"].astype(np.int64) <= np.kron(tile_len, np.ones((16, 16), np.int64))).all()
    for v in grads.values():
        assert np.isfinite(v).all()
    # linearity of the backward in dL/dpix: g(2 dL) == 2 g(dL) exactly (power-of-two scaling)
    _, g2 = helpers.run_cuda(dict(case, dL_dpix=2 * case["dL_dpix"]), want_internals=False)
    helpers.compare_float("linearity colors", g2["colors"], 2 * grads["colors"], rel=1e-5)
    helpers.compare_float("linearity means3D", g2["means3D"], 2 * grads["means3D"], rel=1e-5)


# ------------------------------------------------------------------------------------------------
# fused SplaTAM frame step vs the CPU oracle of the reference's torch code (two passes, autograd)
# ------------------------------------------------------------------------------------------------
def _frame_case(P=3000, W=128, H=96, K=3, seed=7, iso=False, scale_mult=6.0):
    from activesplat_b200 import scenes
    params, cams = scenes.s_room(P, K, seed, W=W, H=H, scale_mult=scale_mult)
    if iso:
        params["log_scales"] = params["log_scales"][:, :1].contiguous()
    rots, trans = scenes.camera_track(cams)
    rots = rots * 1.7          # un-normalised on purpose
    params["cam_unnorm_rots"], params["cam_trans"] = rots, trans
    g = torch.Generator().manual_seed(seed + 1)
    gt_rgb = torch.rand(3, H, W, generator=g)
    gt_depth = torch.rand(1, H, W, generator=g) * 6
    gt_depth[gt_depth < 0.6] = 0.0  # invalid-depth pixels exercise the mask
    return params, cams, gt_rgb, gt_depth


@pytest.mark.parametrize("iso", [False, True])
def test_fused_tracking_get_loss_matches_oracle(iso):
    """splatam.get_loss(tracking=True) (fused path) vs the oracle's restatement of splatam.py:172-301:
    loss components and the camera-pose gradients."""
    from activesplat_b200 import splatam
    from oracle import splatam_oracle as so
    params, cams, gt_rgb, gt_depth = _frame_case(iso=iso)
    W, H, t_idx = cams[0].width, cams[0].height, 1
    lw = dict(im=0.5, depth=1.0)
    # --- oracle (CPU) ---
    pc = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    cam_c = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device="cpu")
    cur_c = {"cam": cam_c, "im": gt_rgb, "depth": gt_depth, "w2c": torch.eye(4)}
    loss_c, wl_c, _ = so.get_loss(pc, cur_c, t_idx, lw, True, 0.5, True, False, tracking=True)
    loss_c.backward()
    # --- fused (GPU) ---
    dev = "cuda:0"
    pg = {k: v.clone().to(dev).requires_grad_(True) for k, v in params.items()}
    cam_g = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device=dev)
    cur_g = {"cam": cam_g, "im": gt_rgb.to(dev), "depth": gt_depth.to(dev), "w2c": torch.eye(4, device=dev)}
    variables = {"max_2D_radius": torch.zeros(pg["means3D"].shape[0], device=dev)}
    loss_g, variables, wl_g = splatam.get_loss(pg, cur_g, variables, t_idx, lw, True, 0.5, True, False, tracking=True)
    loss_g.backward()
    for k in ("loss", "im", "depth"):
        helpers.compare_float("loss " + k, wl_g[k].detach().cpu().numpy(), wl_c[k].detach().numpy())
    helpers.compare_float("grad cam_unnorm_rots", pg["cam_unnorm_rots"].grad.cpu().numpy(),
                          pc["cam_unnorm_rots"].grad.numpy())
    helpers.compare_float("grad cam_trans", pg["cam_trans"].grad.cpu().numpy(), pc["cam_trans"].grad.numpy())
    # tracking detaches means and rotations (splatam.py:178-182); colours/opacities/scales still get gradients
    assert pg["means3D"].grad is None and pg["unnorm_rotations"].grad is None
    assert pc["means3D"].grad is None and pc["unnorm_rotations"].grad is None
    for k in ("rgb_colors", "logit_opacities", "log_scales"):
        helpers.compare_grad("grad " + k, pg[k].grad.cpu().numpy(), pc[k].grad.numpy())
    assert int(variables["seen"].sum()) > 100


@pytest.mark.parametrize("iso", [False, True])
def test_fused_frame_gaussian_gradients_match_oracle(iso):
    """The fused frame backward down to the RAW parameters (sigmoid/exp/normalize/quat_mult/transform
    chain) with both Gaussian and camera gradients on, against torch autograd over the oracle rasterizer."""
    from activesplat_b200 import splatam
    from activesplat_b200.frame import FrameWorkspace, _FusedTrackingLoss
    from oracle import splatam_oracle as so
    params, cams, gt_rgb, gt_depth = _frame_case(P=2500, seed=11, iso=iso)
    W, H, t_idx = cams[0].width, cams[0].height, 2
    lw = dict(im=0.5, depth=1.0)
    pc = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    cam_c = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device="cpu")
    cur_c = {"cam": cam_c, "im": gt_rgb, "depth": gt_depth, "w2c": torch.eye(4)}
    # do_ba=True: gaussians_grad and camera_grad both on; tracking-style sums via a hand-made loss
    tg = so.transform_to_frame(pc, t_idx, gaussians_grad=True, camera_grad=True)
    rv = so.transformed_params2rendervar(pc, tg)
    drv = so.transformed_params2depthplussilhouette(pc, cur_c["w2c"], tg)
    im, radius, _, _ = so.Renderer(cam_c)(**rv)
    ds, _, _, _ = so.Renderer(cam_c)(**drv)
    mask = ((gt_depth > 0) & (ds[1:2] > 0.5)).detach()
    loss_c = 0.5 * torch.abs(gt_rgb - im)[torch.tile(mask, (3, 1, 1))].sum() + torch.abs(gt_depth - ds[0:1])[mask].sum()
    loss_c.backward()
    dev = "cuda:0"
    pg = {k: v.clone().to(dev).requires_grad_(True) for k, v in params.items()}
    cam_g = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device=dev)
    ws = FrameWorkspace(pg["means3D"].shape[0], W, H, dev, capacity=200_000)
    out = _FusedTrackingLoss.apply(pg["means3D"], pg["rgb_colors"], pg["unnorm_rotations"], pg["logit_opacities"],
                                   pg["log_scales"], pg["cam_unnorm_rots"], pg["cam_trans"], ws, cam_g, t_idx,
                                   torch.eye(4, device=dev), gt_rgb.to(dev), gt_depth.to(dev), 0.5, 1.0, 0.5, True, True)
    out[0].backward()
    R, overflow = ws.last_status()
    assert not overflow and R > 0
    helpers.compare_float("loss", out[0].detach().cpu().numpy(), loss_c.detach().numpy())
    helpers.compare_exact("radii", ws.radii.cpu().numpy(), radius.numpy())
    helpers.compare_float("im", ws.color6[:3].cpu().numpy(), im.detach().numpy())
    helpers.compare_float("depth_sil", ws.color6[3:].cpu().numpy(), ds.detach().numpy())
    for k in ("means3D", "rgb_colors", "unnorm_rotations", "logit_opacities", "log_scales", "cam_unnorm_rots", "cam_trans"):
        if iso and k == "unnorm_rotations":
            # equal scales make the covariance rotation-invariant: the analytic gradient is zero and both
            # sides return cancellation noise, so only its magnitude is checked
            assert np.abs(pg[k].grad.cpu().numpy()).max() <= 1e-3 * np.abs(pg["log_scales"].grad.cpu().numpy()).max()
            continue
        helpers.compare_grad("grad " + k, pg[k].grad.cpu().numpy(), pc[k].grad.numpy())
    # capacity overflow is flagged, not written out of bounds
    ws_small = FrameWorkspace(pg["means3D"].shape[0], W, H, dev, capacity=max(R // 3, 16))
    ws_small.forward({k: v.detach() for k, v in pg.items()}, cam_g, t_idx)
    R2, overflow2 = ws_small.last_status()
    assert overflow2 and R2 == R
    assert ws_small.ensure_capacity(int(1.25 * R2))
    ws_small.forward({k: v.detach() for k, v in pg.items()}, cam_g, t_idx)
    assert ws_small.last_status() == (R, False)
    helpers.compare_exact("re-run after growing", ws_small.color6.cpu().numpy(), ws.color6.cpu().numpy())


def test_chunked_forward_equals_the_single_walk():
    """Crowded low-resolution view (GetOpacity camera, ~10^4 records per tile): blending every tile's list in chunks of
    1024 records (pass A transmittance, pass B the reference loop from the true T, combine) gives the images of the
    single serial walk to rounding; the median depth and the 1e-4 stop rule fire in the right chunk."""
    from activesplat_b200 import scenes, splatam
    from activesplat_b200.frame import FrameWorkspace, identity_pose
    dev = torch.device("cuda:0")
    params_cpu, _ = scenes.s_room(700_000, 1, seed=4)
    params = {k: v.to(dev) for k, v in params_cpu.items()}
    rot, tr = identity_pose(dev)
    p = dict(params, cam_unnorm_rots=rot, cam_trans=tr)
    worst_len = 0
    for view in scenes.candidate_views(6, seed=4)[::2]:
        cam, sil_w2c = splatam._camera_one_upload(view.width, view.height, view.k(), view.w2c, 0.01, 100, 1.0, dev)
        ws = FrameWorkspace(700_000, view.width, view.height, dev)
        with torch.no_grad():
            ws.forward(p, cam, 0, sil_w2c=sil_w2c, bg_rgb=torch.ones(3, device=dev))      # calibrates, sees the deep lists
            ws.forward(p, cam, 0, sil_w2c=sil_w2c, bg_rgb=torch.ones(3, device=dev))
            assert not ws._chunked
            ref = [t.clone() for t in (ws.color6, ws.depth, ws.opacity)]
            R, over = ws.last_status()
            assert not over
            ws.forward(p, cam, 0, sil_w2c=sil_w2c, bg_rgb=torch.ones(3, device=dev), forward_only=True)
            assert ws._chunked
            got = [t.clone() for t in (ws.color6, ws.depth, ws.opacity)]
        worst_len = max(worst_len, R // 80)
        helpers.compare_float("chunked colour6", got[0].cpu().numpy(), ref[0].cpu().numpy(), rel=1e-5)
        helpers.compare_float("chunked opacity", got[2].cpu().numpy(), ref[2].cpu().numpy(), rel=1e-5)
        # a pixel whose T lands within rounding of 0.5 may take the neighbouring Gaussian's depth
        helpers.compare_float("chunked median depth", got[1].cpu().numpy(), ref[1].cpu().numpy(), rel=1e-6, flip_bound=1.0)
        with pytest.raises(RuntimeError, match="forward_only"):
            ws.backward(dL_dpix=torch.zeros_like(ws.color6))
    assert worst_len > 2048   # several chunks per tile on average: the split was exercised


def test_workspace_guards():
    """One frame per workspace: the backward of a node whose buffers were overwritten raises (instead of differentiating
    another frame); a second backward raises; a no-grad render through a workspace whose binning buffer is too small
    is repaired; the screen-space gradient the fused path does not keep says so."""
    from activesplat_b200 import frame, splatam
    params, cams, gt_rgb, gt_depth = helpers.frame_case(P=1500, W=96, H=64, seed=5)
    dev = "cuda:0"
    p = {k: v.clone().to(dev).requires_grad_(True) for k, v in params.items()}
    cam = splatam.setup_camera(96, 64, cams[0].k(), np.eye(4), device=dev)
    cur = {"cam": cam, "im": gt_rgb.to(dev), "depth": gt_depth.to(dev), "w2c": torch.eye(4, device=dev)}
    var = {"max_2D_radius": torch.zeros(1500, device=dev)}
    lw = dict(im=0.5, depth=1.0)
    loss0, var, _ = splatam.get_loss(p, cur, var, 0, lw, False, 0.98, True, False, mapping=True)
    loss1, var, _ = splatam.get_loss(p, cur, var, 1, lw, False, 0.98, True, False, mapping=True)
    with pytest.raises(RuntimeError, match="another frame was rendered"):
        loss0.backward()
    loss1.backward(retain_graph=True)
    with pytest.raises(RuntimeError, match="a second time"):
        loss1.backward()
    with pytest.raises(NotImplementedError, match="means2D"):
        var["means2D"].grad
    # a no-grad render does not disturb a frame whose backward is still to come (separate workspace) ...
    loss2, var, _ = splatam.get_loss(p, cur, var, 2, lw, False, 0.98, True, False, mapping=True)
    with torch.no_grad():
        splatam.render_view({k: v.detach() for k, v in p.items()}, cams[0].w2c, cams[0].k(), 96, 64)
    loss2.backward()
    # ... and repairs an overflow of its binning buffer instead of returning a background image
    small = frame.FrameWorkspace(1500, 96, 64, dev, capacity=64)
    with torch.no_grad():
        im_s, _, op_s, _ = splatam.render_view({k: v.detach() for k, v in p.items()}, cams[0].w2c, cams[0].k(), 96, 64, ws=small)
        im_r, _, op_r, _ = splatam.render_view({k: v.detach() for k, v in p.items()}, cams[0].w2c, cams[0].k(), 96, 64)
    assert small.capacity > 64 and float(op_r.max()) > 0.5
    assert torch.equal(im_s, im_r) and torch.equal(op_s, op_r)


def test_fused_adam_matches_torch_adam():
    """as_adam_step vs torch.optim.Adam(eps=1e-15) as configured by initialize_optimizer (splatam.py:118-124)."""
    from activesplat_b200.frame import FusedAdam, PARAM_ORDER, pack_flat
    from activesplat_b200 import scenes
    lrs = dict(means3D=1e-4, rgb_colors=2.5e-3, unnorm_rotations=1e-3, logit_opacities=5e-2, log_scales=1e-3)
    for iso in (False, True):
        params, _ = scenes.s_frustum(5000, 64, 64, 3)
        if iso:
            params["log_scales"] = params["log_scales"][:, :1].contiguous()
        ref = {k: v.clone().requires_grad_(True) for k, v in params.items()}
        opt = torch.optim.Adam([{"params": [ref[k]], "lr": lrs[k]} for k in PARAM_ORDER], lr=0.0, eps=1e-15)
        dev = "cuda:0"
        mine = {k: v.clone().to(dev) for k, v in params.items()}
        fa = FusedAdam(mine, lrs)
        g = torch.Generator().manual_seed(5)
        for step in range(4):
            grads = {k: torch.randn(ref[k].shape, generator=g) * (10.0 ** (step - 2)) for k in PARAM_ORDER}
            for k in PARAM_ORDER:
                ref[k].grad = grads[k].clone()
            opt.step()
            fa.step(pack_flat(grads, 5000).to(dev))   # the flat segment layout of the frame backward
        for k in PARAM_ORDER:
            helpers.compare_float("adam " + k, mine[k].cpu().numpy(), ref[k].detach().numpy(), rel=1e-5)


def test_adam_in_chunks_equals_adam_in_one_launch():
    """as_adam_step_range over [0,a) [a,b) [b,n) with the step advanced by the last chunk == as_adam_step_counted,
    bit for bit, over several steps (dist.allreduce_and_step relies on it)."""
    from activesplat_b200.frame import FusedAdam, PARAM_ORDER
    lrs = dict(means3D=1e-4, rgb_colors=2.5e-3, unnorm_rotations=1e-3, logit_opacities=5e-2, log_scales=1e-3)
    dev = "cuda:0"
    g = torch.Generator().manual_seed(9)
    for P, iso in ((5003, False), (4096, True)):
        shapes = {"means3D": 3, "rgb_colors": 3, "unnorm_rotations": 4, "logit_opacities": 1, "log_scales": 1 if iso else 3}
        a = {k: torch.randn(P, w, generator=g).to(dev) for k, w in shapes.items()}
        b = {k: v.clone() for k, v in a.items()}
        oa, ob = FusedAdam(a, lrs), FusedAdam(b, lrs)
        n = ob.n_flat
        cuts = [0, (n // 3) // 4 * 4, (2 * n // 3) // 4 * 4, n]
        for step in range(3):
            flat = (torch.randn(14 * P, generator=g) * 10.0 ** (step - 1)).to(dev)
            oa.step(flat)
            for i in range(3):
                ob.step_range(flat, cuts[i], cuts[i + 1], advance=(i == 2))
        assert float(oa.step_state[0]) == 3.0 and float(ob.step_state[0]) == 3.0
        for k in PARAM_ORDER:
            assert torch.equal(a[k], b[k]), k
        assert torch.equal(oa.exp_avg[:n], ob.exp_avg[:n]) and torch.equal(oa.exp_avg_sq[:n], ob.exp_avg_sq[:n])


@pytest.mark.parametrize("width", [96, 90])   # 90: not a multiple of 4, the SSIM kernels then stage their tiles with scalar loads
def test_mapping_get_loss_and_adam_steps_match_oracle(width):
    """get_loss(mapping=True) (0.8 L1 + 0.2 (1-SSIM) colour loss, masked mean depth loss, splatam.py:239-249)
    + Adam, three iterations: losses, gradients of all five tensors, and the parameters after the steps,
    against the CPU oracle (two-pass oracle rasterizer + torch autograd + torch.optim.Adam)."""
    from activesplat_b200 import splatam
    from activesplat_b200.frame import FusedAdam, PARAM_ORDER, workspace_for
    from oracle import splatam_oracle as so
    params, cams, gt_rgb, gt_depth = _frame_case(P=2000, W=width, H=80, seed=21)
    W, H, t_idx = cams[0].width, cams[0].height, 1
    lw = dict(im=0.5, depth=1.0)
    lrs = dict(means3D=1e-4, rgb_colors=2.5e-3, unnorm_rotations=1e-3, logit_opacities=5e-2, log_scales=1e-3,
               cam_unnorm_rots=0.0, cam_trans=0.0)
    pc = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    opt_c = torch.optim.Adam([{"params": [v], "name": k, "lr": lrs[k]} for k, v in pc.items()], lr=0.0, eps=1e-15)
    cam_c = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device="cpu")
    cur_c = {"cam": cam_c, "im": gt_rgb, "depth": gt_depth, "w2c": torch.eye(4)}
    dev = "cuda:0"
    pg = {k: v.clone().to(dev).requires_grad_(True) for k, v in params.items()}
    opt_g = FusedAdam(pg, lrs)
    cam_g = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device=dev)
    cur_g = {"cam": cam_g, "im": gt_rgb.to(dev), "depth": gt_depth.to(dev), "w2c": torch.eye(4, device=dev)}
    variables = {"max_2D_radius": torch.zeros(pg["means3D"].shape[0], device=dev)}
    for it in range(3):
        loss_c, wl_c, _ = so.get_loss(pc, cur_c, t_idx, lw, False, 0.98, True, False, mapping=True)
        opt_c.zero_grad(set_to_none=True)
        loss_c.backward()
        loss_g, variables, wl_g = splatam.get_loss(pg, cur_g, variables, t_idx, lw, False, 0.98, True, False, mapping=True)
        for p in pg.values():
            p.grad = None
        loss_g.backward()
        for k in ("loss", "im", "depth"):
            helpers.compare_float(f"it{it} loss {k}", wl_g[k].detach().cpu().numpy(), wl_c[k].detach().numpy())
        if it == 0:
            for k in PARAM_ORDER:
                helpers.compare_grad("grad " + k, pg[k].grad.cpu().numpy(), pc[k].grad.numpy())
            assert pg["cam_trans"].grad is None and pc["cam_trans"].grad is None  # camera_grad=False in mapping
        opt_c.step()
        opt_g.step(workspace_for(pg, cam_g).grad_params)
    for k in PARAM_ORDER:
        d_ref = (pc[k].detach() - params[k]).numpy()
        d_got = (pg[k].detach().cpu() - params[k]).numpy()
        # Adam's first steps are sign-like (|update| ~ lr): compare the parameter DRIFT, elementwise
        helpers.compare_float("drift " + k, d_got, d_ref, rel=2e-3, flip_fraction=5e-3, flip_bound=2.5)


def test_render_view_matches_oracle_render():
    """render_view (GetOpacity per-view body) vs the oracle's get_rendervars + render (two passes, white bg)."""
    from activesplat_b200 import scenes, splatam
    from oracle import raster_oracle as ro
    params, _ = scenes.s_room(4000, 3, 2, scale_mult=12.0)
    cam = scenes.candidate_views(3, seed=2)[2]
    dev = "cuda:0"
    pg = {k: v.to(dev) for k, v in params.items()}
    im, depth, opacity, sil = splatam.render_view(pg, cam.w2c, cam.k(), cam.width, cam.height)
    a = scenes.activated(params)
    view, proj, campos = cam.matrices()
    feats = helpers.depth_sil_features(a["means3D"], cam.w2c).numpy()
    common = (a["opacities"].numpy(), a["scales"].numpy(), a["rotations"].numpy(), 1.0, None, view, proj, cam.tanfovx,
              cam.tanfovy, cam.height, cam.width)
    st1 = ro.forward(a["means3D"].numpy(), a["colors_precomp"].numpy(), *common, np.ones(3, np.float32))
    st2 = ro.forward(a["means3D"].numpy(), feats, *common, np.zeros(3, np.float32))
    helpers.compare_float("im", im.cpu().numpy(), st1["color"])
    helpers.compare_float("median depth", depth.cpu().numpy(), st1["depth"])
    helpers.compare_float("opacity", opacity.cpu().numpy(), st1["opacity"])
    helpers.compare_float("silhouette", sil.cpu().numpy(), st2["color"][1:2])


def test_fused_frame_edge_cases():
    """Fused frame path: image sizes that are not multiples of 16, P not a multiple of 256, a frame in which
    every Gaussian is culled, and the NaN-loss signal on binning overflow."""
    from activesplat_b200 import scenes, splatam
    from activesplat_b200.frame import FrameWorkspace, fused_tracking_loss
    from oracle import splatam_oracle as so
    dev = "cuda:0"
    params, cams = scenes.s_room(1777, 2, 31, W=150, H=100, scale_mult=5.0)
    rots, trans = scenes.camera_track(cams)
    params["cam_unnorm_rots"], params["cam_trans"] = rots, trans
    g = torch.Generator().manual_seed(3)
    gt_rgb, gt_depth = torch.rand(3, 100, 150, generator=g), torch.rand(1, 100, 150, generator=g) * 5
    lw = dict(im=0.5, depth=1.0)
    pc = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    cam_c = splatam.setup_camera(150, 100, cams[0].k(), np.eye(4), device="cpu")
    loss_c, wl_c, aux = so.get_loss(pc, {"cam": cam_c, "im": gt_rgb, "depth": gt_depth, "w2c": torch.eye(4)}, 1, lw,
                                    True, 0.9, True, False, tracking=True)
    loss_c.backward()
    pg = {k: v.clone().to(dev).requires_grad_(True) for k, v in params.items()}
    cam_g = splatam.setup_camera(150, 100, cams[0].k(), np.eye(4), device=dev)
    cur_g = {"cam": cam_g, "im": gt_rgb.to(dev), "depth": gt_depth.to(dev), "w2c": torch.eye(4, device=dev)}
    variables = {"max_2D_radius": torch.zeros(1777, device=dev)}
    loss_g, variables, _ = splatam.get_loss(pg, cur_g, variables, 1, lw, True, 0.9, True, False, tracking=True)
    loss_g.backward()
    helpers.compare_float("loss", loss_g.detach().cpu().numpy(), loss_c.detach().numpy())
    helpers.compare_float("grad cam_trans", pg["cam_trans"].grad.cpu().numpy(), pc["cam_trans"].grad.numpy())
    helpers.compare_exact("seen", variables["seen"].cpu().numpy(), (aux["radius"] > 0).numpy())
    helpers.compare_exact("max_2D_radius", variables["max_2D_radius"].cpu().numpy(), aux["radius"].float().numpy())
    # every Gaussian behind the camera: background image, zero gradients, loss = the GT sums under the mask
    far = {k: v.detach().clone() for k, v in pg.items()}
    far["means3D"] = far["means3D"] * 0 + torch.tensor([0.0, 0.0, -50.0], device=dev)
    far["cam_unnorm_rots"] = torch.tensor([1.0, 0, 0, 0], device=dev).reshape(1, 4, 1).repeat(1, 1, 2).contiguous()
    far["cam_trans"] = torch.zeros(1, 3, 2, device=dev)
    ws = FrameWorkspace(1777, 150, 100, dev, capacity=4096)
    ws.forward(far, cam_g, 0)
    assert ws.last_status() == (0, False) and int(ws.radii.max()) == 0
    assert float(ws.color6.abs().max()) == 0.0 and bool((ws.depth == 15.0).all()) and float(ws.opacity.max()) == 0.0
    ws.tracking_loss(gt_rgb.to(dev), gt_depth.to(dev), 0.5, 1.0, 0.9, False, False)
    gp, gc = ws.backward(gaussians_grad=True, camera_grad=True)
    assert float(gp.abs().max()) == 0.0 and float(gc.abs().max()) == 0.0
    # overflow: the loss turns NaN instead of silently using a truncated instance list
    small = FrameWorkspace(1777, 150, 100, dev, capacity=64)
    out, _ = fused_tracking_loss({k: v.detach() for k, v in pg.items()}, cam_g, 1, None, gt_rgb.to(dev),
                                 gt_depth.to(dev), 0.5, 1.0, 0.9, True, True, ws=small)
    assert bool(torch.isnan(out).all()) and small.last_status()[1]


def _load_reference_build():
    """The UNMODIFIED reference rasterizer built by oracle/build_ref.py (travels to the GPU box); None if absent."""
    import importlib.util
    path = os.path.join(helpers.ROOT, "oracle", "_ref", "ref_dgr_C.so")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location("ref_dgr_C", path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


@pytest.mark.parametrize("size", ["cfg1_100k_640x480", "room_1M_lookaround"])
def test_full_size_matches_reference_build_live(size):
    """Full BASELINE sizes against the reference CUDA build running on the same GPU (no stored vectors needed):
    radii, sorted keys, sorted instance list, tile ranges and n_contrib bit-identical; images and gradients 1e-4."""
    ref = _load_reference_build()
    if ref is None:
        pytest.skip("oracle/_ref/ref_dgr_C.so not built (needs /root/reference at build time)")
    sys_path_added = os.path.join(helpers.ROOT, "oracle")
    import importlib
    mg = importlib.import_module("oracle.make_golden")
    from activesplat_b200 import scenes
    if size == "cfg1_100k_640x480":
        params, cam = scenes.s_frustum(100_000, 640, 480, 0)
    else:
        params, _ = scenes.s_room(1_000_000, 1, 4)
        cam = scenes.candidate_views(3, seed=4)[1]
    a = scenes.activated(params)
    view, proj, campos = cam.matrices()
    case = dict(means3D=a["means3D"].numpy(), colors=a["colors_precomp"].numpy(), opacities=a["opacities"].numpy(),
                scales=a["scales"].numpy(), rotations=a["rotations"].numpy(), scale_modifier=1.0, view=view, proj=proj,
                campos=campos, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, H=cam.height, W=cam.width,
                bg=np.array([0.3, 0.2, 0.1], np.float32),
                dL_dpix=np.random.default_rng(100).uniform(-1, 1, (3, cam.height, cam.width)).astype(np.float32))
    r = mg.run_reference(ref, case)
    out, grads = helpers.run_cuda(case)
    assert out["num_rendered"] == int(r["num_rendered"])
    for k in ("radii", "tiles_touched", "keys", "point_list", "ranges", "n_contrib"):
        helpers.compare_exact(k, out[k], r[k])
    vis = r["radii"] > 0
    helpers.compare_exact("means2D", out["means2D"][vis], r["means2D"][vis])
    helpers.compare_exact("depths", out["depths"][vis], r["depths"][vis])
    for k in ("color", "depth", "opacity", "final_T"):
        helpers.compare_float(k, out[k], r[k])
    for k in GRADS:
        # the reference sums ~10^2-10^4 atomics per Gaussian in arbitrary order: compare at 1e-4 of the scale,
        # allowing a small fraction of large-cancellation entries
        helpers.compare_float("grad_" + k, grads[k], r["grad_" + k], rel=1e-4, flip_fraction=2e-3, flip_bound=5e-2)


def test_add_new_gaussians_matches_oracle():
    """Silhouette-driven densification (splatam.py:332-379), SURVEY 8f rank 1: same new Gaussians as the oracle."""
    from activesplat_b200 import splatam
    from oracle import splatam_oracle as so
    params, cams, gt_rgb, gt_depth = _frame_case(P=1500, W=96, H=80, seed=41, scale_mult=2.0)  # map with holes
    gt_depth = gt_depth * 0.6 + 0.4
    gt_depth[0, 30:50, 20:60] = 0.5   # a new foreground object in front of the mapped surface
    gt_depth[0, :5, :] = 0.0          # invalid depth
    W, H, t_idx = cams[0].width, cams[0].height, 1
    k = torch.tensor(cams[0].k()).float()
    cam_c = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device="cpu")
    cur_c = {"cam": cam_c, "im": gt_rgb, "depth": gt_depth, "w2c": torch.eye(4), "intrinsics": k}
    mask_c, new_c = so.add_new_gaussians({k_: v.clone() for k_, v in params.items()}, cur_c, 0.5, t_idx, "anisotropic")
    dev = "cuda:0"
    pg = {k_: torch.nn.Parameter(v.clone().to(dev)) for k_, v in params.items()}
    cam_g = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device=dev)
    cur_g = {"cam": cam_g, "im": gt_rgb.to(dev), "depth": gt_depth.to(dev), "w2c": torch.eye(4, device=dev),
             "intrinsics": k.to(dev)}
    variables = {"timestep": torch.zeros(1500, device=dev), "max_2D_radius": torch.zeros(1500, device=dev)}
    P0 = 1500
    pg, variables = splatam.add_new_gaussians(pg, variables, cur_g, 0.5, t_idx, "projective", "anisotropic")
    n_new = pg["means3D"].shape[0] - P0
    assert new_c is not None and n_new > 100
    # a pixel whose silhouette sits within rounding of the threshold may flip; everything else must agree
    assert abs(n_new - new_c["means3D"].shape[0]) <= max(2, int(1e-3 * n_new))
    if n_new == new_c["means3D"].shape[0]:
        for k_ in ("means3D", "rgb_colors", "unnorm_rotations", "logit_opacities", "log_scales"):
            helpers.compare_float("new " + k_, pg[k_][P0:].detach().cpu().numpy(), new_c[k_].numpy(), rel=1e-5)
            helpers.compare_exact("old " + k_, pg[k_][:P0].detach().cpu().numpy(), params[k_].numpy())
    assert variables["timestep"].shape[0] == P0 + n_new and float(variables["timestep"][P0:].min()) == t_idx
    assert variables["max_2D_radius"].shape[0] == P0 + n_new and pg["means3D"].requires_grad


def test_noncontiguous_rotations_use_the_forward_values():
    """SplaTAM's anisotropic path feeds the rasterizer column-major rotations.  The stock reference backward reads
    that memory as if it were row-major (rasterize_points.cu:173,175 vs :102,:104) and returns scale/rotation
    gradients of scrambled quaternions; we differentiate w.r.t. the values the forward used (== oracle, == the
    reference itself when it is given contiguous inputs)."""
    from activesplat_b200 import _C
    from activesplat_b200.rasterizer import GaussianRasterizationSettings, make_rasterize_function
    case = helpers.make_case("frustum_small")
    dev = "cuda:0"
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    e = torch.Tensor([])
    rot_cm = t(case["rotations"].T.copy()).T          # same values, column-major memory, like quat_mult's output
    assert not rot_cm.is_contiguous()

    def run(backend, rot):
        rs = GaussianRasterizationSettings(case["H"], case["W"], case["tanfovx"], case["tanfovy"], t(case["bg"]), 1.0,
                                           t(case["view"]).unsqueeze(0), t(case["proj"]).unsqueeze(0), 0, t(case["campos"]),
                                           False, False)
        leaves = dict(means3D=t(case["means3D"]).requires_grad_(True), scales=t(case["scales"]).requires_grad_(True),
                      rot=rot.clone(memory_format=torch.preserve_format).requires_grad_(True))
        color, _, _, _ = make_rasterize_function(backend).apply(leaves["means3D"], torch.zeros_like(leaves["means3D"]), e,
                                                                t(case["colors"]), t(case["opacities"]), leaves["scales"],
                                                                leaves["rot"], e, rs)
        (color * t(case["dL_dpix"])).sum().backward()
        return {k: v.grad.contiguous().cpu().numpy() for k, v in leaves.items()}

    st, g = helpers.run_oracle(case)
    ours = run(_C, rot_cm)
    helpers.compare_grad("grad scales vs oracle", ours["scales"], g["scales"])
    helpers.compare_grad("grad rotations vs oracle", ours["rot"], g["rotations"])
    ref = _load_reference_build()
    if ref is None:
        pytest.skip("reference build not present")
    theirs_cm = run(ref, rot_cm)
    theirs_c = run(ref, rot_cm.contiguous())
    helpers.compare_grad("grad means3D agrees with the stock reference", ours["means3D"], theirs_cm["means3D"])
    helpers.compare_grad("grad rotations == reference given contiguous rotations", ours["rot"], theirs_c["rot"])
    helpers.compare_grad("grad scales == reference given contiguous rotations", ours["scales"], theirs_c["scales"])
    # and the stock reference on the column-major tensor is far off (documented divergence)
    assert np.abs(theirs_cm["rot"] - g["rotations"]).max() > 0.05 * np.abs(g["rotations"]).max()


def test_tile_buckets_and_global_sort_agree_at_full_size():
    """cfg2-sized frame (500k Gaussians, 640x480) through both binning modes of the fused path: the sorted instance
    list, the tile ranges, radii and every output image must be IDENTICAL (same list => same arithmetic), and each
    tile's list must be ordered by (depth bits, Gaussian id) -- the order the reference's stable radix sort of
    (tile, depth) keys yields (rasterizer_impl.cu:305-310).  Size-independent properties at BASELINE's full size."""
    import ctypes
    from activesplat_b200 import capi, scenes, splatam
    from activesplat_b200.frame import FrameWorkspace
    dev = torch.device("cuda:0")
    P, W, H, K = 500_000, 640, 480, 2
    params_cpu, cams = scenes.s_room(P, K, seed=1, W=W, H=H)
    rots, trans = scenes.camera_track(cams)
    params = {k: v.to(dev) for k, v in params_cpu.items()}
    params["cam_unnorm_rots"], params["cam_trans"] = rots.to(dev), trans.to(dev)
    cam = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device=dev)
    cap = 1_000_000
    L = capi.lib()
    lay = capi.AsScratchLayout()
    capi.check(L.as_scratch_layout(P, W, H, cap, 6, ctypes.byref(lay)))
    tiles = ((W + 15) // 16) * ((H + 15) // 16)

    def run(mode, t):
        ws = FrameWorkspace(P, W, H, dev, capacity=cap, binning=mode)
        ws.forward(params, cam, t)
        R, overflow = ws.last_status()
        assert not overflow and 0 < R < cap
        point_list = ws.binning[lay.point_list: lay.point_list + 4 * R].view(torch.int32).cpu().numpy()
        ranges = ws.img[lay.ranges: lay.ranges + 8 * tiles].view(torch.int32).cpu().numpy().reshape(tiles, 2)
        depths = ws.geom[lay.depths: lay.depths + 4 * P].view(torch.int32).cpu().numpy()
        out = dict(R=R, point_list=point_list, ranges=ranges, depths=depths, radii=ws.radii.cpu().numpy(),
                   color=ws.color6.cpu().numpy(), depth=ws.depth.cpu().numpy(), opacity=ws.opacity.cpu().numpy())
        return ws, out

    for t in range(K):
        ws0, a = run(0, t)
        ws1, b = run(1, t)
        assert a["R"] == b["R"]
        for k in ("radii", "ranges", "point_list"):
            helpers.compare_exact(f"{k} (global sort vs tile buckets)", b[k], a[k])
        for k in ("color", "depth", "opacity"):
            helpers.compare_exact(f"{k} image (global sort vs tile buckets)", b[k].view(np.int32), a[k].view(np.int32))
        # sortedness: inside every tile the key (depth bits, id) strictly increases
        key = (a["depths"][a["point_list"]].astype(np.int64) << 32) | a["point_list"].astype(np.int64)
        inc = np.diff(key) > 0
        starts = a["ranges"][:, 0][a["ranges"][:, 1] > a["ranges"][:, 0]]
        boundary = np.zeros(a["R"] - 1, bool)
        boundary[starts[starts > 0] - 1] = True          # the pair straddling two tiles is exempt
        assert (inc | boundary).all()
        # the ranges tile the list without gaps or overlaps, in tile order
        nonempty = a["ranges"][a["ranges"][:, 1] > a["ranges"][:, 0]]
        assert nonempty[0, 0] == 0 and nonempty[-1, 1] == a["R"] and (nonempty[1:, 0] == nonempty[:-1, 1]).all()
        assert int(ws1.status[3]) == int((a["ranges"][:, 1] - a["ranges"][:, 0]).max())   # largest bucket reported


@pytest.mark.parametrize("shape", [(640, 480, 300_000), (320, 240, 150_000), (192, 160, 80_000), (64, 48, 20_000)])
def test_blend_pipelines_repeat_exactly(shape):
    """The blend kernels' warps are not held together by a block barrier per batch (the last warp out of a stage
    refills it; the forward stops through a data-less barrier completion).  A race there would show up as a run that
    differs from the others: the forward of one frame, repeated, must give bit-identical images, final T and
    n_contrib, and the backward the same gradients up to the order of its float reductions.  The shapes cover every
    CTA shape (8, 4, 2 and 1 warps per CTA) and lists of many batches per tile."""
    from activesplat_b200 import scenes, splatam
    from activesplat_b200.frame import FrameWorkspace
    W, H, P = shape
    dev = torch.device("cuda:0")
    params_cpu, cams = scenes.s_room(P, 2, seed=7, W=W, H=H)
    rots, trans = scenes.camera_track(cams)
    params = {k: v.to(dev) for k, v in params_cpu.items()}
    params["cam_unnorm_rots"], params["cam_trans"] = rots.to(dev), trans.to(dev)
    cam = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device=dev)
    ws = FrameWorkspace(P, W, H, dev, capacity=8 * P, binning=1)
    dL = torch.from_numpy(np.random.default_rng(5).uniform(-1, 1, (6, H, W)).astype(np.float32)).to(dev)
    first = None
    for rep in range(12):
        ws.forward(params, cam, 1)
        R, overflow = ws.last_status()
        assert not overflow and R > 0
        imgs = (ws.color6.clone(), ws.depth.clone(), ws.opacity.clone())
        gp, _ = ws.backward(dL_dpix=dL, gaussians_grad=True, camera_grad=False)
        gp = gp[:14 * P].clone()
        assert torch.isfinite(gp).all()
        if first is None:
            first = (imgs, gp)
            continue
        for a, b in zip(imgs, first[0]):
            assert torch.equal(a, b), f"forward run {rep} differs from run 0 at {shape}"
        scale = float(first[1].abs().max())
        assert float((gp - first[1]).abs().max()) <= 2e-5 * scale, f"backward run {rep} differs from run 0 at {shape}"


def test_tile_buckets_larger_than_shared_memory_still_sort_exactly():
    """A tiny image under a big map: every tile's bucket exceeds what one CTA sorts in shared memory, so the
    in-place (global-memory) network runs.  Same sorted list and images as the global radix sort."""
    import ctypes
    from activesplat_b200 import capi, scenes, splatam
    from activesplat_b200.frame import FrameWorkspace
    dev = torch.device("cuda:0")
    P, W, H = 60_000, 48, 40
    params_cpu, cams = scenes.s_room(P, 1, seed=21, W=W, H=H)
    rots, trans = scenes.camera_track(cams)
    params = {k: v.to(dev) for k, v in params_cpu.items()}
    params["cam_unnorm_rots"], params["cam_trans"] = rots.to(dev), trans.to(dev)
    cam = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device=dev)
    cap = 400_000
    L = capi.lib()
    lay = capi.AsScratchLayout()
    capi.check(L.as_scratch_layout(P, W, H, cap, 6, ctypes.byref(lay)))
    outs = []
    for mode in (0, 1):
        ws = FrameWorkspace(P, W, H, dev, capacity=cap, binning=mode)
        ws.forward(params, cam, 0)
        R, overflow = ws.last_status()
        assert not overflow
        outs.append((R, ws.binning[lay.point_list: lay.point_list + 4 * R].view(torch.int32).cpu().numpy(),
                     ws.color6.cpu().numpy().view(np.int32), ws.opacity.cpu().numpy().view(np.int32),
                     int(ws.status[3])))
    assert outs[1][4] > L.as_tile_sort_capacity(), "the case must exercise the big-bucket path"
    assert outs[0][0] == outs[1][0]
    helpers.compare_exact("point_list", outs[1][1], outs[0][1])
    helpers.compare_exact("colour", outs[1][2], outs[0][2])
    helpers.compare_exact("opacity", outs[1][3], outs[0][3])


def test_score_candidate_views_equals_per_view_renders():
    """The low-host-cost sweep (dist.score_candidate_views) renders exactly what splatam.render_view renders."""
    from activesplat_b200 import dist as asd, scenes, splatam
    dev = torch.device("cuda:0")
    params_cpu, _ = scenes.s_room(4000, 1, seed=31, scale_mult=4.0)
    params = {k: v.to(dev) for k, v in params_cpu.items()}
    views = scenes.candidate_views(7, seed=31)
    scores, ops = asd.score_candidate_views(params, views, n_streams=3, return_opacity=True)
    assert scores.shape == (7, 2)
    for j, c in enumerate(views):
        _, _, opacity, _ = splatam.render_view(params, c.w2c, c.k(), c.width, c.height)
        helpers.compare_exact(f"opacity of view {j}", ops[j].cpu().numpy().view(np.int32),
                              opacity[0].cpu().numpy().view(np.int32))
        assert abs(float(scores[j, 0]) - float((1.0 - opacity).sum())) <= 1e-4 * max(1.0, float(scores[j, 0]))
    asd._VIEW_WORKSPACES.clear()


def test_auto_binning_falls_back_to_the_global_sort_without_synchronising():
    """A workspace calibrated on a sparse frame that later sees crowded tiles notices it from the asynchronously
    copied status words and switches to the global sort for the following frames; results never change."""
    from activesplat_b200 import capi, scenes, splatam
    from activesplat_b200.frame import FrameWorkspace
    dev = torch.device("cuda:0")
    P, W, H = 60_000, 48, 40
    params_cpu, cams = scenes.s_room(P, 1, seed=21, W=W, H=H)
    rots, trans = scenes.camera_track(cams)
    params = {k: v.to(dev) for k, v in params_cpu.items()}
    params["cam_unnorm_rots"], params["cam_trans"] = rots.to(dev), trans.to(dev)
    cam = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device=dev)
    sparse = {k: (v[:500].contiguous() if k not in ("cam_unnorm_rots", "cam_trans") else v) for k, v in params.items()}
    ws = FrameWorkspace(P, W, H, dev)            # auto
    ws.P = 500
    ws.forward(sparse, cam, 0)                   # calibration frame: few Gaussians, small buckets
    assert ws.binning_mode == 1
    ws.P = P
    ws.ensure_capacity(400_000)
    ref = FrameWorkspace(P, W, H, dev, capacity=400_000, binning=0)
    ref.forward(params, cam, 0)
    for _ in range(4):                           # crowded frames: correct from the first one, global sort soon after
        ws.forward(params, cam, 0)
        torch.cuda.synchronize()
        helpers.compare_exact("colour", ws.color6.cpu().numpy().view(np.int32), ref.color6.cpu().numpy().view(np.int32))
    assert ws.binning_mode == 0


def test_peer_sharded_adam_single_rank_equals_fused_adam():
    """as_adam_step_peer with one rank (its own gradient, the whole map as its slab) against as_adam_step: same
    arithmetic per element, so the parameters agree to the last bit after several steps."""
    from activesplat_b200 import dist as asd
    from activesplat_b200.frame import FusedAdam, PARAM_ORDER
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(5)
    P = 3001
    shapes = {"means3D": 3, "rgb_colors": 3, "unnorm_rotations": 4, "logit_opacities": 1, "log_scales": 3}
    lrs = {"means3D": 1e-4, "rgb_colors": 2.5e-3, "unnorm_rotations": 1e-3, "logit_opacities": 5e-2, "log_scales": 1e-3}
    a = {k: torch.randn(P, w, generator=g).to(dev) for k, w in shapes.items()}
    b = {k: v.clone() for k, v in a.items()}
    flat = torch.zeros((P, 14), device=dev)
    opt_a = FusedAdam(a, lrs)
    opt_b = asd.PeerShardedAdam(b, lrs, flat)
    for step in range(4):
        flat.copy_(torch.randn(P, 14, generator=g) * 10.0 ** (step - 2))
        opt_a.step(flat)
        opt_b.step()
    for k in PARAM_ORDER:
        helpers.compare_float("peer adam " + k, b[k].cpu().numpy(), a[k].cpu().numpy(), rel=1e-6)
