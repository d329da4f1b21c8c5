"""Fused SplaTAM frame step on top of the C ABI (as_frame_forward / as_frame_backward /
as_tracking_loss / as_adam_step).

`FrameWorkspace` owns the persistent device buffers of one (P, W, H) configuration so that a step
allocates nothing and never synchronises with the host; `fused_tracking_loss` is the autograd entry
that `splatam.get_loss(tracking=True)` uses; `FusedAdam` mirrors `torch.optim.Adam` as configured
by `initialize_optimizer` (src/mapper/splatam/splatam.py:118-124) over the five Gaussian tensors.
"""
import ctypes

import torch

from . import capi

PARAM_ORDER = ("means3D", "rgb_colors", "unnorm_rotations", "logit_opacities", "log_scales")
SLOTS = {"means3D": (0, 3), "rgb_colors": (3, 6), "unnorm_rotations": (6, 10), "logit_opacities": (10, 11),
         "log_scales": (11, 14)}


class FrameWorkspace:
    """Persistent scratch + outputs for the fused frame path.

    The binning buffer holds `capacity` instances.  The number actually needed is only known on the
    device; it comes back with the losses (`last_status()` reads it).  If a frame overflowed, its
    outputs are invalid: `ensure_capacity()` grows the buffer and the caller re-runs the frame.
    One workspace supports one forward/backward pair in flight.
    """

    def __init__(self, P, width, height, device, capacity=None, want_means2D_grad=False, binning=None):
        """binning: None = choose on the first frame (tile buckets unless a tile holds more instances than one
        CTA sorts in shared memory), 0 = global radix sort, 1 = tile buckets.  Results are identical."""
        L = capi.lib()
        self.binning_mode = 1 if binning is None else int(binning)
        self._binning_auto = binning is None
        self._status_host = None     # pinned copy of `status` (auto binning keeps an eye on the largest bucket)
        self._status_event = None
        self.P, self.W, self.H, self.device = int(P), int(width), int(height), torch.device(device)
        self.capacity = int(capacity) if capacity else max(4 * self.P, 1 << 16)
        with torch.cuda.device(self.device):
            u8 = dict(dtype=torch.uint8, device=self.device)
            f32 = dict(dtype=torch.float32, device=self.device)
            self.geom = torch.zeros(L.as_geometry_bytes(self.P), **u8)   # accumulators must start at zero
            self.img = torch.empty(L.as_image_bytes(self.W, self.H), **u8)
            self.binning = torch.empty(L.as_binning_bytes(self.capacity, 6), **u8)
            self.status = torch.zeros(4, dtype=torch.int32, device=self.device)
            self.color6 = torch.empty((6, self.H, self.W), **f32)
            self.depth = torch.empty((1, self.H, self.W), **f32)
            self.opacity = torch.empty((1, self.H, self.W), **f32)
            self.radii = torch.empty((self.P,), dtype=torch.int32, device=self.device)
            self.dL = torch.empty((6, self.H, self.W), **f32)
            self.losses = torch.zeros(8, **f32)
            self.loss_scratch = torch.empty(L.as_loss_scratch_floats(self.W, self.H), **f32)
            self.grad_params = torch.zeros((self.P, 14), **f32)
            self.grad_means2D = torch.zeros((self.P, 3), **f32) if want_means2D_grad else None
            self.grad_cam = torch.zeros(8, **f32)
            self.pose_scratch = torch.empty(L.as_pose_scratch_floats(self.P), **f32)
            self.bg6 = torch.zeros(6, **f32)
        self._args = None
        self._keep = None
        self._bg_src = None
        self.map_scratch = None
        self.calibrated = capacity is not None  # an explicit capacity is taken as is (and so is the binning mode)

    # -- buffer management ---------------------------------------------------------------------
    def ensure_capacity(self, needed):
        if needed <= self.capacity:
            return False
        self.capacity = int(needed) + 1024
        with torch.cuda.device(self.device):
            self.binning = torch.empty(capi.lib().as_binning_bytes(self.capacity, 6), dtype=torch.uint8,
                                       device=self.device)
        return True

    def last_status(self):
        """(num_rendered, overflowed) of the last frame -- synchronises."""
        s = self.status.tolist()
        return int(s[0]), bool(s[1])

    # -- forward / backward --------------------------------------------------------------------
    def forward(self, params, cam, time_idx, sil_w2c=None, bg_rgb=None, debug=False, max_2D_radius=None, seen=None):
        """as_frame_forward.  params: dict with the five Gaussian tensors + cam_unnorm_rots [1,4,T] +
        cam_trans [1,3,T]; cam: GaussianRasterizationSettings.  Results land in color6/depth/opacity/radii."""
        L = capi.lib()
        a = capi.AsFrameArgs()
        keep = []

        def ptr(t):
            if t.dtype != torch.float32 or not t.is_contiguous():
                t = t.contiguous().float()
            keep.append(t)
            return t.data_ptr()

        a.P, a.width, a.height = self.P, self.W, self.H
        a.isotropic = 1 if params["log_scales"].shape[1] == 1 else 0
        a.means3D = ptr(params["means3D"].detach())
        a.rgb_colors = ptr(params["rgb_colors"].detach())
        a.unnorm_rotations = ptr(params["unnorm_rotations"].detach())
        a.logit_opacities = ptr(params["logit_opacities"].detach())
        a.log_scales = ptr(params["log_scales"].detach())
        rots, trans = params["cam_unnorm_rots"].detach(), params["cam_trans"].detach()
        T = rots.shape[-1]
        assert rots.is_contiguous() and trans.is_contiguous() and rots.dtype == torch.float32
        a.cam_unnorm_rot = rots.data_ptr() + 4 * int(time_idx)
        a.cam_rot_stride = T
        a.cam_trans = trans.data_ptr() + 4 * int(time_idx)
        a.cam_trans_stride = T
        keep += [rots, trans]
        a.sil_w2c = ptr(sil_w2c) if sil_w2c is not None else None
        a.viewmatrix = ptr(cam.viewmatrix)
        a.projmatrix = ptr(cam.projmatrix)
        bg = cam.bg if bg_rgb is None else bg_rgb
        if bg is not self._bg_src:  # refresh the 6-channel background only when the caller's object changes
            self.bg6[:3].copy_(torch.as_tensor(bg, dtype=torch.float32, device=self.device))
            self._bg_src = bg
        a.background = self.bg6.data_ptr()
        a.tan_fovx, a.tan_fovy, a.scale_modifier = float(cam.tanfovx), float(cam.tanfovy), float(cam.scale_modifier)
        a.geom_buffer, a.image_buffer, a.binning_buffer = self.geom.data_ptr(), self.img.data_ptr(), self.binning.data_ptr()
        a.binning_capacity = self.capacity
        a.status = self.status.data_ptr()
        a.out_color, a.out_depth, a.out_opacity = self.color6.data_ptr(), self.depth.data_ptr(), self.opacity.data_ptr()
        a.radii = self.radii.data_ptr()
        a.debug = int(bool(debug))
        a.max_2D_radius = max_2D_radius.data_ptr() if max_2D_radius is not None else None
        a.seen = seen.data_ptr() if seen is not None else None
        capturing = torch.cuda.is_current_stream_capturing()
        if self._binning_auto and self.binning_mode == 1 and self._status_event is not None and not capturing:
            # the status words of an EARLIER frame, copied back asynchronously: if a tile's bucket outgrew what a CTA
            # sorts in shared memory, later frames use the global sort (no synchronisation, one or two frames late;
            # the oversized frames themselves were still sorted correctly, only slowly)
            if self._status_event.query():
                if int(self._status_host[3]) > L.as_tile_sort_capacity():
                    self.binning_mode = 0
                self._status_event = None
        a.binning_mode = self.binning_mode
        with torch.cuda.device(self.device):
            stream = torch.cuda.current_stream(self.device).cuda_stream
            capi.check(L.as_frame_forward(ctypes.byref(a), stream))
            if self._binning_auto and self.binning_mode == 1 and self.calibrated and self._status_event is None \
                    and not capturing:
                if self._status_host is None:
                    self._status_host = torch.zeros(4, dtype=torch.int32).pin_memory()
                self._status_host.copy_(self.status, non_blocking=True)
                self._status_event = torch.cuda.Event()
                self._status_event.record(torch.cuda.current_stream(self.device))
            if not self.calibrated and not torch.cuda.is_current_stream_capturing():
                # first frame of this workspace: ONE synchronising look at the instance count sizes the binning
                # buffer with head-room (2x); after that the path never synchronises and an overflow shows up as
                # a NaN loss / last_status()
                self.calibrated = True
                s = self.status.tolist()
                rerun = self.ensure_capacity(2 * int(s[0]))
                if self._binning_auto and self.binning_mode == 1 and (s[1] or int(s[3]) > L.as_tile_sort_capacity()):
                    # crowded tiles (low-resolution wide-angle views): the global sort handles any distribution
                    self.binning_mode = a.binning_mode = 0
                    rerun = True
                if rerun:
                    a.binning_buffer, a.binning_capacity = self.binning.data_ptr(), self.capacity
                    capi.check(L.as_frame_forward(ctypes.byref(a), stream))
        self._args, self._keep = a, keep
        return self.color6, self.radii, self.depth, self.opacity

    def tracking_loss(self, gt_rgb, gt_depth, w_im, w_depth, sil_thres, use_sil_for_loss, mask_color):
        L = capi.lib()
        keep = self._keep

        def ptr(t):
            if t.dtype != torch.float32 or not t.is_contiguous():
                t = t.contiguous().float()
            keep.append(t)
            return t.data_ptr()

        with torch.cuda.device(self.device):
            capi.check(L.as_tracking_loss(self.W, self.H, self.color6.data_ptr(), ptr(gt_rgb), ptr(gt_depth),
                                          float(w_im), float(w_depth), float(sil_thres), int(bool(use_sil_for_loss)),
                                          int(bool(mask_color)), self.dL.data_ptr(), self.losses.data_ptr(),
                                          self.loss_scratch.data_ptr(), self.status.data_ptr(),
                                          torch.cuda.current_stream(self.device).cuda_stream))
        return self.losses

    def mapping_loss(self, gt_rgb, gt_depth, w_im, w_depth):
        """as_mapping_loss on the last forward's colour6: fills self.dL and self.losses."""
        L = capi.lib()
        keep = self._keep

        def ptr(t):
            if t.dtype != torch.float32 or not t.is_contiguous():
                t = t.contiguous().float()
            keep.append(t)
            return t.data_ptr()

        with torch.cuda.device(self.device):
            if self.map_scratch is None:
                self.map_scratch = torch.empty(L.as_mapping_loss_scratch_floats(self.W, self.H), dtype=torch.float32,
                                               device=self.device)
            capi.check(L.as_mapping_loss(self.W, self.H, self.color6.data_ptr(), ptr(gt_rgb), ptr(gt_depth),
                                         float(w_im), float(w_depth), self.dL.data_ptr(), self.losses.data_ptr(),
                                         self.map_scratch.data_ptr(), self.status.data_ptr(),
                                         torch.cuda.current_stream(self.device).cuda_stream))
        return self.losses

    def backward(self, dL_dpix=None, grad_scale=None, gaussians_grad=True, camera_grad=False, accumulate=False,
                 grad_cam_out=None, grad_cam_stride=1):
        """as_frame_backward for the last forward.  Returns (grad_params [P,14], grad_cam).
        grad_cam_out/grad_cam_stride: write the 7 pose gradients at grad_cam_out[k * stride] instead."""
        L = capi.lib()
        b = capi.AsFrameBackwardArgs()
        b.frame = ctypes.pointer(self._args)
        dL = self.dL if dL_dpix is None else dL_dpix.contiguous()
        b.dL_dpix = dL.data_ptr()
        b.grad_scale = grad_scale.data_ptr() if grad_scale is not None else None
        b.gaussians_grad, b.camera_grad = int(bool(gaussians_grad)), int(bool(camera_grad))
        b.grad_params = self.grad_params.data_ptr()
        b.grad_means2D = self.grad_means2D.data_ptr() if self.grad_means2D is not None else None
        gc = self.grad_cam if grad_cam_out is None else grad_cam_out
        b.grad_cam = gc.data_ptr()
        b.grad_cam_stride = int(grad_cam_stride)
        b.pose_scratch = self.pose_scratch.data_ptr()
        b.accumulate = int(bool(accumulate))
        self._keep.append(dL)
        with torch.cuda.device(self.device):
            capi.check(L.as_frame_backward(ctypes.byref(b), torch.cuda.current_stream(self.device).cuda_stream))
        return self.grad_params, gc


_WORKSPACES = {}


def workspace_for(params, cam):
    """One cached workspace per (P, W, H, device)."""
    key = (int(params["means3D"].shape[0]), int(cam.image_width), int(cam.image_height),
           params["means3D"].device.index)
    ws = _WORKSPACES.get(key)
    if ws is None:
        ws = FrameWorkspace(key[0], key[1], key[2], params["means3D"].device)
        _WORKSPACES[key] = ws
    return ws


class _FusedTrackingLoss(torch.autograd.Function):
    """The frame rendered from the raw parameters and its loss as ONE autograd node.
    use_sil is a bool  -> tracking loss: w_im * sum_mask|gt_rgb - im| + w_depth * sum_mask|gt_depth - depth|
                          (splatam.py:230-247);
    use_sil == "mapping" -> mapping loss: w_im * (0.8 L1 + 0.2 (1 - SSIM)) + w_depth * masked mean depth L1
                          (splatam.py:239-249)."""

    @staticmethod
    def forward(ctx, means3D, rgb, rots, logit_op, log_scales, cam_rots, cam_trans, ws, cam, time_idx, sil_w2c,
                gt_rgb, gt_depth, w_im, w_depth, sil_thres, use_sil, mask_color, max_2D_radius=None, seen=None):
        params = {"means3D": means3D, "rgb_colors": rgb, "unnorm_rotations": rots, "logit_opacities": logit_op,
                  "log_scales": log_scales, "cam_unnorm_rots": cam_rots, "cam_trans": cam_trans}
        ws.forward(params, cam, time_idx, sil_w2c=sil_w2c, max_2D_radius=max_2D_radius, seen=seen)
        if isinstance(use_sil, str):
            ws.mapping_loss(gt_rgb, gt_depth, w_im, w_depth)
        else:
            ws.tracking_loss(gt_rgb, gt_depth, w_im, w_depth, sil_thres, use_sil, mask_color)
        ctx.ws, ctx.time_idx = ws, int(time_idx)
        ctx.shapes = (cam_rots.shape, cam_trans.shape, log_scales.shape[1])
        ctx.gaussians_grad = any(ctx.needs_input_grad[:5])
        ctx.camera_grad = any(ctx.needs_input_grad[5:7])
        ctx.need = tuple(ctx.needs_input_grad[:7])
        return ws.losses[:3].clone()

    @staticmethod
    def backward(ctx, g):
        ws = ctx.ws
        # d(total)/d... ; the im/depth components are returned for logging only
        scale = g[0:1].contiguous()
        T = ctx.shapes[0][-1]
        gcam = torch.zeros((7, T), dtype=torch.float32, device=g.device) if ctx.camera_grad else None
        gp, gc = ws.backward(grad_scale=scale, gaussians_grad=ctx.gaussians_grad, camera_grad=ctx.camera_grad,
                             grad_cam_out=gcam[:, ctx.time_idx:] if gcam is not None else None, grad_cam_stride=T)
        outs = [None] * 20
        if ctx.gaussians_grad:
            ls_w = ctx.shapes[2]
            views = (gp[:, 0:3], gp[:, 3:6], gp[:, 6:10], gp[:, 10:11], gp[:, 11:11 + ls_w])
            for i in range(5):
                if ctx.need[i]:
                    outs[i] = views[i]
        if ctx.camera_grad:
            outs[5], outs[6] = gcam[0:4].unsqueeze(0), gcam[4:7].unsqueeze(0)
        return tuple(outs)


class _FusedFrameRender(torch.autograd.Function):
    """The frame render as one differentiable node: raw parameters + pose -> colour6 [6,H,W]
    (rgb | depth, silhouette, depth^2).  Replaces transform_to_frame + the rendervar builders + the two
    Renderer calls of get_loss (splatam.py:191-212); median depth / opacity / radii are side outputs."""

    @staticmethod
    def forward(ctx, means3D, rgb, rots, logit_op, log_scales, cam_rots, cam_trans, ws, cam, time_idx, sil_w2c, bg_rgb):
        params = {"means3D": means3D, "rgb_colors": rgb, "unnorm_rotations": rots, "logit_opacities": logit_op,
                  "log_scales": log_scales, "cam_unnorm_rots": cam_rots, "cam_trans": cam_trans}
        ws.forward(params, cam, time_idx, sil_w2c=sil_w2c, bg_rgb=bg_rgb)
        ctx.ws, ctx.time_idx = ws, int(time_idx)
        ctx.shapes = (cam_rots.shape, cam_trans.shape, log_scales.shape[1])
        ctx.need = tuple(ctx.needs_input_grad[:7])
        out = (ws.color6.clone(), ws.radii, ws.depth, ws.opacity)
        ctx.mark_non_differentiable(*out[1:])
        return out

    @staticmethod
    def backward(ctx, g_color6, _r, _d, _o):
        ws = ctx.ws
        gaussians_grad, camera_grad = any(ctx.need[:5]), any(ctx.need[5:7])
        T = ctx.shapes[0][-1]
        gcam = torch.zeros((7, T), dtype=torch.float32, device=g_color6.device) if camera_grad else None
        gp, gc = ws.backward(dL_dpix=g_color6, gaussians_grad=gaussians_grad, camera_grad=camera_grad,
                             grad_cam_out=gcam[:, ctx.time_idx:] if gcam is not None else None, grad_cam_stride=T)
        outs = [None] * 12
        if gaussians_grad:
            views = (gp[:, 0:3], gp[:, 3:6], gp[:, 6:10], gp[:, 10:11], gp[:, 11:11 + ctx.shapes[2]])
            for i in range(5):
                if ctx.need[i]:
                    outs[i] = views[i]
        if camera_grad:
            outs[5], outs[6] = gcam[0:4].unsqueeze(0), gcam[4:7].unsqueeze(0)
        return tuple(outs)


def fused_render(params, cam, time_idx, sil_w2c=None, bg_rgb=None, ws=None):
    """(colour6 [6,H,W], radii, median depth, opacity, workspace); colour6 is differentiable w.r.t. `params`."""
    ws = ws or workspace_for(params, cam)
    c6, radii, depth, opacity = _FusedFrameRender.apply(
        params["means3D"], params["rgb_colors"], params["unnorm_rotations"], params["logit_opacities"],
        params["log_scales"], params["cam_unnorm_rots"], params["cam_trans"], ws, cam, time_idx, sil_w2c, bg_rgb)
    return c6, radii, depth, opacity, ws


_IDENTITY_POSE = {}


def identity_pose(device):
    """cam_unnorm_rots [1,4,1] = (1,0,0,0), cam_trans [1,3,1] = 0: Gaussians stay in the world frame and
    the real w2c lives in the raster settings (render() / GetOpacity)."""
    key = torch.device(device)
    if key not in _IDENTITY_POSE:
        r = torch.zeros((1, 4, 1), device=device)
        r[0, 0, 0] = 1.0
        _IDENTITY_POSE[key] = (r, torch.zeros((1, 3, 1), device=device))
    return _IDENTITY_POSE[key]


def fused_tracking_loss(params, cam, time_idx, sil_w2c, gt_rgb, gt_depth, w_im, w_depth, sil_thres,
                        use_sil_for_loss, mask_color, ws=None, max_2D_radius=None, seen=None):
    """Returns (losses[3] = weighted total / im / depth as a device tensor, workspace)."""
    ws = ws or workspace_for(params, cam)
    out = _FusedTrackingLoss.apply(params["means3D"], params["rgb_colors"], params["unnorm_rotations"],
                                   params["logit_opacities"], params["log_scales"], params["cam_unnorm_rots"],
                                   params["cam_trans"], ws, cam, time_idx, sil_w2c, gt_rgb, gt_depth, w_im, w_depth,
                                   sil_thres, use_sil_for_loss, mask_color, max_2D_radius, seen)
    return out, ws


def fused_mapping_loss(params, cam, time_idx, sil_w2c, gt_rgb, gt_depth, w_im, w_depth, ws=None, max_2D_radius=None,
                       seen=None):
    """Mapping-loss twin of fused_tracking_loss: (losses[3] = weighted total / im / depth, workspace)."""
    ws = ws or workspace_for(params, cam)
    out = _FusedTrackingLoss.apply(params["means3D"], params["rgb_colors"], params["unnorm_rotations"],
                                   params["logit_opacities"], params["log_scales"], params["cam_unnorm_rots"],
                                   params["cam_trans"], ws, cam, time_idx, sil_w2c, gt_rgb, gt_depth, w_im, w_depth,
                                   0.0, "mapping", False, max_2D_radius, seen)
    return out, ws


class FusedAdam:
    """torch.optim.Adam(lr per tensor, betas (0.9, 0.999), eps) over the five Gaussian tensors in one kernel.
    State is flat [P,14]; `step(grad_flat)` consumes the flat gradient buffer of the frame backward."""

    def __init__(self, params, lrs, betas=(0.9, 0.999), eps=1e-15):
        self.params = params
        self.lrs = (ctypes.c_float * 5)(*[float(lrs[k]) for k in PARAM_ORDER])
        self.betas, self.eps, self.t = betas, eps, 0
        P, dev = params["means3D"].shape[0], params["means3D"].device
        self.exp_avg = torch.zeros((P, 14), dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros((P, 14), dtype=torch.float32, device=dev)

    @torch.no_grad()
    def step(self, grad_flat):
        L = capi.lib()
        self.t += 1
        p = self.params
        dev = p["means3D"].device
        for k in PARAM_ORDER:
            assert p[k].is_contiguous() and p[k].dtype == torch.float32
        with torch.cuda.device(dev):
            capi.check(L.as_adam_step(int(p["means3D"].shape[0]), 1 if p["log_scales"].shape[1] == 1 else 0,
                                      p["means3D"].data_ptr(), p["rgb_colors"].data_ptr(),
                                      p["unnorm_rotations"].data_ptr(), p["logit_opacities"].data_ptr(),
                                      p["log_scales"].data_ptr(), grad_flat.data_ptr(), self.exp_avg.data_ptr(),
                                      self.exp_avg_sq.data_ptr(), self.lrs, self.betas[0], self.betas[1], self.eps,
                                      self.t, torch.cuda.current_stream(dev).cuda_stream))


class GraphedStep:
    """Replays a fixed step (e.g. `get_loss(...)` + `backward()` for one keyframe) as ONE CUDA graph launch.

    The fused frame path never synchronises with the host and allocates nothing of its own, so the ~25
    kernels of a step can be captured once and replayed; that removes the launch gaps between the many
    short kernels (scan, sort passes, gather, loss).  `fn` must read its inputs from tensors that stay alive
    and in place (update them with copy_ before a replay); its outputs are overwritten by every replay.
    """

    def __init__(self, fn, warmup=2):
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(warmup):
                fn()
        cur.wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.out = fn()

    def __call__(self):
        self.graph.replay()
        return self.out
