"""Fused SplaTAM frame step on top of the C ABI (as_frame_forward / as_frame_backward /
as_tracking_loss / as_adam_step).

`FrameWorkspace` owns the persistent device buffers of one (P, W, H) configuration so that a step
allocates nothing and never synchronises with the host; `fused_tracking_loss` is the autograd entry
that `splatam.get_loss(tracking=True)` uses; `FusedAdam` mirrors `torch.optim.Adam` as configured
by `initialize_optimizer` (src/mapper/splatam/splatam.py:118-124) over the five Gaussian tensors.
"""
import ctypes
import warnings

import torch

from . import capi

PARAM_ORDER = ("means3D", "rgb_colors", "unnorm_rotations", "logit_opacities", "log_scales")
# The flat gradient / Adam-state layout (include/activesplat_b200.h, AsFrameBackwardArgs.grad_params): 14 P floats, the
# five tensors one after the other, each in its parameter's own [P,w] layout.  SEGMENTS: name -> (offset / P, width).
SEGMENTS = {"means3D": (0, 3), "rgb_colors": (3, 3), "unnorm_rotations": (6, 4), "logit_opacities": (10, 1),
            "log_scales": (11, 3)}


def flat_views(flat, P, scale_width=3):
    """The five per-tensor views [P,w] of a flat buffer in the segment layout (contiguous, no copies)."""
    flat = flat.reshape(-1)
    out = {}
    for k in PARAM_ORDER:
        off, w = SEGMENTS[k]
        w = scale_width if k == "log_scales" else w
        out[k] = flat[off * P: off * P + w * P].view(P, w)
    return out


def pack_flat(tensors, P, device=None):
    """A flat buffer in the segment layout from per-tensor [P,w] arrays (tests, tools)."""
    dev = device if device is not None else tensors["means3D"].device
    flat = torch.zeros(14 * P, dtype=torch.float32, device=dev)
    for k, v in flat_views(flat, P, tensors["log_scales"].shape[1]).items():
        v.copy_(tensors[k])
    return flat


class FrameWorkspace:
    """Persistent scratch + outputs for the fused frame path.

    The binning buffer holds `capacity` instances.  The number actually needed is only known on the
    device; it comes back with the losses (`last_status()` reads it).  If a frame overflowed, its
    outputs are invalid: `ensure_capacity()` grows the buffer and the caller re-runs the frame.
    One workspace supports one forward/backward pair in flight.
    """

    def __init__(self, P, width, height, device, capacity=None, want_means2D_grad=False, binning=None):
        """binning: None = choose on the first frame (tile buckets unless a tile holds more instances than one
        CTA sorts in shared memory), 0 = global radix sort, 1 = tile buckets.  Results are identical.
        P is the number of Gaussians of the first frame; `resize()` follows a growing map (densification)."""
        L = capi.lib()
        self.binning_mode = 1 if binning is None else int(binning)
        self._binning_auto = binning is None
        self._status_host = None     # pinned copy of `status` of an earlier frame, fetched without synchronising
        self._status_event = None
        self.W, self.H, self.device = int(width), int(height), torch.device(device)
        self.capacity = int(capacity) if capacity else max(4 * int(P), 1 << 16)
        self._want_means2D_grad = bool(want_means2D_grad)
        self.P = self.cap_P = 0
        with torch.cuda.device(self.device):
            u8 = dict(dtype=torch.uint8, device=self.device)
            f32 = dict(dtype=torch.float32, device=self.device)
            self.img = torch.empty(L.as_image_bytes(self.W, self.H), **u8)
            self.binning = torch.empty(L.as_binning_bytes(self.capacity, 6), **u8)
            self.status = torch.zeros(4, dtype=torch.int32, device=self.device)
            self.color6 = torch.empty((6, self.H, self.W), **f32)
            self.depth = torch.empty((1, self.H, self.W), **f32)
            self.opacity = torch.empty((1, self.H, self.W), **f32)
            self.dL = torch.empty((6, self.H, self.W), **f32)
            self.losses = torch.zeros(8, **f32)
            self.loss_scratch = torch.empty(L.as_loss_scratch_floats(self.W, self.H), **f32)
            self.grad_cam = torch.zeros(8, **f32)
            self.bg6 = torch.zeros(6, **f32)
        self.resize(int(P))
        self._args = None
        self._keep = None
        self._bg_src = None
        self.map_scratch = None
        self.calibrated = capacity is not None  # an explicit capacity is taken as is (and so is the binning mode)
        self._calib_P = None
        self.generation = 0          # bumped by every forward: a backward must belong to the latest one
        self.overflow_seen = None    # (instances, capacity) of an earlier frame that did not fit, until reported
        self._graph_pinned = False   # a CUDA graph captured this workspace's buffers: never free them
        self._retired = []
        self._side = None            # side stream for clearing the gradient buffer next to the forward
        self._grad_cleared = False
        self._last_R = 0             # instance count of an earlier frame (asynchronous status copy)
        self._chunked = False        # the last forward blended in chunks (no n_contrib: no backward may follow)
        self._chunk_scratch = None

    # -- buffer management ---------------------------------------------------------------------
    def resize(self, P):
        """Follow the map to P Gaussians.  The per-Gaussian buffers are allocated for cap_P >= P and grow
        geometrically (x1.5), so a map that gains a few thousand Gaussians per frame reallocates rarely and the
        binning buffer, its calibration and the image-sized buffers are kept."""
        P = int(P)
        if P == self.P:
            return
        L = capi.lib()
        if getattr(self, "_calib_P", None) and not (0.5 * self._calib_P <= P <= 1.25 * self._calib_P):
            self.calibrated = False   # a different map, not a growing one: size the binning buffer again
        with torch.cuda.device(self.device):
            if P > self.cap_P:
                if getattr(self, "_graph_pinned", False):
                    self._retired.append((self._geom, self._radii, self._grad_params, self._grad_means2D, self.pose_scratch))
                self.cap_P = max(P, int(1.5 * self.cap_P))
                u8 = dict(dtype=torch.uint8, device=self.device)
                f32 = dict(dtype=torch.float32, device=self.device)
                self._geom = torch.empty(L.as_geometry_bytes(self.cap_P), **u8)
                self._radii = torch.empty((self.cap_P,), dtype=torch.int32, device=self.device)
                self._grad_params = torch.zeros((self.cap_P * 14,), **f32)
                self._grad_means2D = torch.zeros((self.cap_P * 3,), **f32) if self._want_means2D_grad else None
                self.pose_scratch = torch.empty(L.as_pose_scratch_floats(self.cap_P), **f32)
            # the scratch layout depends on P: the backward accumulators (which the kernels expect to find zeroed, and
            # leave zeroed) land somewhere else, so the whole geometry buffer is cleared (stream-ordered, no sync)
            self._geom.zero_()
            self.P = P
            self.geom = self._geom[: L.as_geometry_bytes(P)]
            self.radii = self._radii[:P]
            self.grad_params = self._grad_params[: P * 14]   # flat segment layout (see SEGMENTS)
            self.grad_means2D = self._grad_means2D[: P * 3].view(P, 3) if self._grad_means2D is not None else None

    def ensure_capacity(self, needed):
        if needed <= self.capacity:
            return False
        if self._graph_pinned:
            self._retired.append(self.binning)
        self.capacity = int(needed) + 1024
        with torch.cuda.device(self.device):
            self.binning = torch.empty(capi.lib().as_binning_bytes(self.capacity, 6), dtype=torch.uint8,
                                       device=self.device)
        return True

    def _poll_status(self):
        """Look at the status words of an EARLIER frame if their asynchronous copy has landed (never waits):
        switches crowded views to the global sort, grows the binning buffer ahead of need (at half full), and
        remembers an overflow so that the caller hears about it."""
        if self._status_event is None or not self._status_event.query():
            return
        self._status_event = None
        n_inst, overflowed, _, largest = (int(x) for x in self._status_host)
        self._last_R = n_inst
        if self._binning_auto and self.binning_mode == 1 and largest > capi.lib().as_tile_sort_capacity():
            self.binning_mode = 0
        if overflowed:
            self.overflow_seen = (n_inst, self._status_cap)
        if 2 * n_inst > self.capacity:
            self.ensure_capacity(2 * n_inst)

    def last_status(self):
        """(num_rendered, overflowed) of the last frame -- synchronises."""
        s = self.status.tolist()
        return int(s[0]), bool(s[1])

    # -- forward / backward --------------------------------------------------------------------
    def forward(self, params, cam, time_idx, sil_w2c=None, bg_rgb=None, debug=False, max_2D_radius=None, seen=None,
                clear_gradient=False, forward_only=False):
        """as_frame_forward.  params: dict with the five Gaussian tensors + cam_unnorm_rots [1,4,T] +
        cam_trans [1,3,T]; cam: GaussianRasterizationSettings.  Results land in color6/depth/opacity/radii.
        clear_gradient: the caller will run a non-accumulating backward with Gaussian gradients on this frame; the flat
        gradient buffer is then cleared now, on a side stream next to the (latency-bound) forward kernels, instead of
        in front of the backward's last kernel.
        forward_only: no backward will follow.  If this workspace's earlier frames had deep lists (>= 2048 records per
        tile on average: low-resolution wide-angle candidate views), the tiles' lists are then blended in chunks, one CTA
        per chunk (AsFrameArgs.chunk_scratch), instead of one serial walk per pixel strip."""
        L = capi.lib()
        self._grad_cleared = False
        # worth it when the lists are deep: two passes over every record buy parallelism along the list, which a frame
        # with a few hundred records per tile does not need (640x480: 0.38 -> 0.33 ms per view without it)
        tiles = ((self.W + 15) // 16) * ((self.H + 15) // 16)
        self._chunked = bool(forward_only and self._binning_auto and self._last_R >= 2048 * tiles)
        self.resize(params["means3D"].shape[0])
        if clear_gradient:
            cur = torch.cuda.current_stream(self.device)
            if self._side is None:
                self._side = torch.cuda.Stream(device=self.device)
            self._side.wait_stream(cur)          # everything that read the buffer so far (Adam, all-reduce) is ahead
            with torch.cuda.stream(self._side):
                self.grad_params.zero_()
            self._grad_cleared = True
        a = capi.AsFrameArgs()
        keep = []

        def ptr(t):
            if t.dtype != torch.float32 or not t.is_contiguous():
                t = t.contiguous().float()
            keep.append(t)
            return t.data_ptr()

        self.resize(params["means3D"].shape[0])
        self.generation += 1
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
        if not capturing:
            # the status words of an EARLIER frame, copied back asynchronously (no synchronisation, one or two frames
            # late): a tile bucket that outgrew what a CTA sorts in shared memory sends later frames to the global sort
            # (the oversized frames themselves were still sorted correctly, only slowly); an instance count beyond half
            # of the binning buffer grows it before it can overflow
            self._poll_status()
            a.binning_buffer, a.binning_capacity = self.binning.data_ptr(), self.capacity
            if self.overflow_seen is not None:
                n_inst, cap = self.overflow_seen
                self.overflow_seen = None
                warnings.warn(f"activesplat_b200: an earlier frame needed {n_inst} Gaussian-tile instances but the binning "
                              f"buffer held {cap}; that frame rendered as background (NaN loss, zero gradients).  The "
                              "buffer has been grown.", RuntimeWarning, stacklevel=3)
        else:
            self._graph_pinned = True
        a.binning_mode = self.binning_mode
        a.chunk_scratch = None
        if self._chunked:
            need = L.as_chunk_scratch_bytes(self.capacity, self.W, self.H)
            if self._chunk_scratch is None or self._chunk_scratch.numel() < need:
                if self._graph_pinned and self._chunk_scratch is not None:
                    self._retired.append(self._chunk_scratch)
                with torch.cuda.device(self.device):
                    self._chunk_scratch = torch.empty(need, dtype=torch.uint8, device=self.device)
            a.chunk_scratch = self._chunk_scratch.data_ptr()
        with torch.cuda.device(self.device):
            stream = torch.cuda.current_stream(self.device).cuda_stream
            capi.check(L.as_frame_forward(ctypes.byref(a), stream))
            if self.calibrated and self._status_event is None and not capturing:
                if self._status_host is None:
                    self._status_host = torch.zeros(4, dtype=torch.int32).pin_memory()
                self._status_host.copy_(self.status, non_blocking=True)
                self._status_cap = self.capacity
                self._status_event = torch.cuda.Event()
                self._status_event.record(torch.cuda.current_stream(self.device))
            if not self.calibrated and not capturing:
                # first frame of this workspace: ONE synchronising look at the instance count sizes the binning
                # buffer with head-room (2x); after that the path never synchronises and an overflow shows up as
                # a NaN loss / last_status()
                self.calibrated, self._calib_P = True, self.P
                s = self.status.tolist()
                self._last_R = int(s[0])
                rerun = self.ensure_capacity(2 * int(s[0]))
                if self._binning_auto and self.binning_mode == 1 and (s[1] or int(s[3]) > L.as_tile_sort_capacity()):
                    # crowded tiles (low-resolution wide-angle views): the global sort handles any distribution
                    self.binning_mode = a.binning_mode = 0
                    rerun = True
                if rerun:
                    a.chunk_scratch = None   # sized for the old capacity; this one frame takes the single walk
                    self._chunked = False
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
        """as_frame_backward for the last forward.  Returns (grad_params: 14 P floats in the flat segment layout, grad_cam).
        grad_cam_out/grad_cam_stride: write the 7 pose gradients at grad_cam_out[k * stride] instead."""
        L = capi.lib()
        if self._chunked:
            raise RuntimeError("activesplat_b200: the last forward of this workspace was forward_only (chunked blend); "
                               "it has no backward")
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
        mode = int(bool(accumulate))
        if self._grad_cleared and gaussians_grad and not accumulate:
            # the buffer was zeroed during the forward (side stream): the backward only writes the visible rows
            torch.cuda.current_stream(self.device).wait_stream(self._side)
            mode = 2
        self._grad_cleared = False
        b.accumulate = mode
        self._keep.append(dL)
        with torch.cuda.device(self.device):
            capi.check(L.as_frame_backward(ctypes.byref(b), torch.cuda.current_stream(self.device).cuda_stream))
        return self.grad_params, gc


_WORKSPACES = {}


def workspace_for(params, cam, role="train"):
    """The cached workspace of (image size, device, role).  The number of Gaussians is not part of the key: a
    workspace follows a growing map (FrameWorkspace.resize), so densification neither leaks the previous size's
    buffers nor re-calibrates.  role "train" serves the differentiable paths (get_loss); no-grad renders
    (render_view, add_new_gaussians) use role "render" so that they cannot overwrite a frame whose backward is
    still to come."""
    dev = params["means3D"].device
    key = (int(cam.image_width), int(cam.image_height), dev.index, role)
    ws = _WORKSPACES.get(key)
    if ws is None:
        ws = FrameWorkspace(int(params["means3D"].shape[0]), key[0], key[1], dev)
        _WORKSPACES[key] = ws
    return ws


_PARAM_GRADS = [True]


class gradients_stay_in_workspace:
    """Context manager for training loops that step with FusedAdam: inside it, backward() of the fused nodes leaves the
    five Gaussian tensors' `.grad` at None and delivers their gradients ONLY in the workspace's flat buffer
    (`workspace.grad_params`, which FusedAdam.step consumes) -- autograd otherwise copies 56 bytes per Gaussian out of
    that buffer into five `.grad` tensors every iteration.  Camera-pose gradients are delivered as usual."""

    def __enter__(self):
        self.prev = _PARAM_GRADS[0]
        _PARAM_GRADS[0] = False
        return self

    def __exit__(self, *a):
        _PARAM_GRADS[0] = self.prev


def _claim_backward(ctx):
    """A workspace holds ONE frame: the backward of a node must find the buffers of its own forward."""
    if getattr(ctx, "spent", False):
        raise RuntimeError("activesplat_b200: backward through a fused frame node a second time; its workspace "
                           "buffers were consumed by the first (run the forward again)")
    if ctx.ws.generation != ctx.generation:
        raise RuntimeError("activesplat_b200: another frame was rendered through the same workspace between this "
                           "node's forward and its backward, so its buffers are gone.  Call backward() before the "
                           "next get_loss / render on this image size, or give each frame in flight its own "
                           "FrameWorkspace (the `ws=` argument of frame.fused_render / fused_tracking_loss).")
    ctx.spent = True


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
        ws.forward(params, cam, time_idx, sil_w2c=sil_w2c, max_2D_radius=max_2D_radius, seen=seen,
                   clear_gradient=any(ctx.needs_input_grad[:5]))
        if isinstance(use_sil, str):
            ws.mapping_loss(gt_rgb, gt_depth, w_im, w_depth)
        else:
            ws.tracking_loss(gt_rgb, gt_depth, w_im, w_depth, sil_thres, use_sil, mask_color)
        ctx.ws, ctx.time_idx, ctx.generation = ws, int(time_idx), ws.generation
        ctx.shapes = (cam_rots.shape, cam_trans.shape, log_scales.shape[1])
        ctx.gaussians_grad = any(ctx.needs_input_grad[:5])
        ctx.camera_grad = any(ctx.needs_input_grad[5:7])
        ctx.need = tuple(ctx.needs_input_grad[:7])
        return ws.losses[:3].clone()

    @staticmethod
    def backward(ctx, g):
        ws = ctx.ws
        _claim_backward(ctx)
        # d(total)/d... ; the im/depth components are returned for logging only
        scale = g[0:1].contiguous()
        T = ctx.shapes[0][-1]
        gcam = torch.zeros((7, T), dtype=torch.float32, device=g.device) if ctx.camera_grad else None
        gp, gc = ws.backward(grad_scale=scale, gaussians_grad=ctx.gaussians_grad, camera_grad=ctx.camera_grad,
                             grad_cam_out=gcam[:, ctx.time_idx:] if gcam is not None else None, grad_cam_stride=T)
        outs = [None] * 20
        if ctx.gaussians_grad:
            # copies: the workspace buffer is rewritten by the next frame, and autograd would otherwise adopt these
            # contiguous views as .grad (use FusedAdam.step(ws.grad_params) to consume the buffer without copies)
            views = flat_views(gp, ws.P, ctx.shapes[2])
            for i, k in enumerate(PARAM_ORDER):
                if ctx.need[i]:
                    outs[i] = views[k].clone() if _PARAM_GRADS[0] else None
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
        ws.forward(params, cam, time_idx, sil_w2c=sil_w2c, bg_rgb=bg_rgb, clear_gradient=any(ctx.needs_input_grad[:5]),
                   forward_only=not any(ctx.needs_input_grad[:7]))
        ctx.ws, ctx.time_idx, ctx.generation = ws, int(time_idx), ws.generation
        ctx.shapes = (cam_rots.shape, cam_trans.shape, log_scales.shape[1])
        ctx.need = tuple(ctx.needs_input_grad[:7])
        # fresh tensors, like the reference's rasterizer returns: the workspace is overwritten by its next frame
        out = (ws.color6.clone(), ws.radii.clone(), ws.depth.clone(), ws.opacity.clone())
        ctx.mark_non_differentiable(*out[1:])
        return out

    @staticmethod
    def backward(ctx, g_color6, _r, _d, _o):
        ws = ctx.ws
        _claim_backward(ctx)
        gaussians_grad, camera_grad = any(ctx.need[:5]), any(ctx.need[5:7])
        T = ctx.shapes[0][-1]
        gcam = torch.zeros((7, T), dtype=torch.float32, device=g_color6.device) if camera_grad else None
        gp, gc = ws.backward(dL_dpix=g_color6, gaussians_grad=gaussians_grad, camera_grad=camera_grad,
                             grad_cam_out=gcam[:, ctx.time_idx:] if gcam is not None else None, grad_cam_stride=T)
        outs = [None] * 12
        if gaussians_grad:
            views = flat_views(gp, ws.P, ctx.shapes[2])
            for i, k in enumerate(PARAM_ORDER):
                if ctx.need[i]:
                    outs[i] = views[k].clone() if _PARAM_GRADS[0] else None
        if camera_grad:
            outs[5], outs[6] = gcam[0:4].unsqueeze(0), gcam[4:7].unsqueeze(0)
        return tuple(outs)


def fused_render(params, cam, time_idx, sil_w2c=None, bg_rgb=None, ws=None):
    """(colour6 [6,H,W], radii, median depth, opacity, workspace); colour6 is differentiable w.r.t. `params`.
    Without a workspace argument, a call that needs no gradient uses the "render" workspace."""
    needs_grad = torch.is_grad_enabled() and any(
        params[k].requires_grad for k in PARAM_ORDER + ("cam_unnorm_rots", "cam_trans"))
    if ws is None:
        ws = workspace_for(params, cam, "train" if needs_grad else "render")
    args = (params["means3D"], params["rgb_colors"], params["unnorm_rotations"], params["logit_opacities"],
            params["log_scales"], params["cam_unnorm_rots"], params["cam_trans"], ws, cam, time_idx, sil_w2c, bg_rgb)
    if needs_grad or torch.cuda.is_current_stream_capturing():
        return (*_FusedFrameRender.apply(*args), ws)
    # a render nobody differentiates: worth one look at the status words (synchronises, as the callers of these paths do
    # anyway) so that a binning overflow is repaired instead of returned as a background image
    for _ in range(3):
        out = _FusedFrameRender.apply(*args)
        n_inst, overflowed = ws.last_status()
        if not overflowed:
            return (*out, ws)
        ws.ensure_capacity(n_inst + n_inst // 4)
    raise RuntimeError(f"activesplat_b200: binning buffer still too small after growing it to {ws.capacity}")


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
    State is flat (14 P floats, segment layout); `step(grad_flat)` consumes the flat gradient buffer of the frame
    backward (`workspace.grad_params`); `step()` without it packs the tensors' `.grad` first."""

    def __init__(self, params, lrs, betas=(0.9, 0.999), eps=1e-15, device_step=True):
        """device_step: keep the step count (and the bias corrections derived from it) in device memory
        (as_adam_step_counted), which makes every step the same launch -- a mapping iteration can then be captured in
        a CUDA graph.  False passes the host-side count to as_adam_step."""
        self.params = params
        self.lrs = (ctypes.c_float * 5)(*[float(lrs[k]) for k in PARAM_ORDER])
        self.betas, self.eps, self.t = betas, eps, 0
        P, dev = params["means3D"].shape[0], params["means3D"].device
        self.exp_avg = torch.zeros(14 * P, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(14 * P, dtype=torch.float32, device=dev)
        self.step_state = torch.zeros(8, dtype=torch.float32, device=dev) if device_step else None

    @torch.no_grad()
    def step(self, grad_flat=None):
        L = capi.lib()
        self.t += 1
        p = self.params
        dev = p["means3D"].device
        if grad_flat is None:
            grad_flat = pack_flat({k: (p[k].grad if p[k].grad is not None else torch.zeros_like(p[k]))
                                   for k in PARAM_ORDER}, int(p["means3D"].shape[0]))
        if grad_flat.numel() < 14 * int(p["means3D"].shape[0]):
            raise ValueError("FusedAdam.step: the flat gradient buffer holds fewer than 14 * P floats")
        for k in PARAM_ORDER:
            assert p[k].is_contiguous() and p[k].dtype == torch.float32
        args = (int(p["means3D"].shape[0]), 1 if p["log_scales"].shape[1] == 1 else 0, p["means3D"].data_ptr(),
                p["rgb_colors"].data_ptr(), p["unnorm_rotations"].data_ptr(), p["logit_opacities"].data_ptr(),
                p["log_scales"].data_ptr(), grad_flat.data_ptr(), self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(),
                self.lrs, self.betas[0], self.betas[1], self.eps)
        with torch.cuda.device(dev):
            stream = torch.cuda.current_stream(dev).cuda_stream
            if self.step_state is not None:
                capi.check(L.as_adam_step_counted(*args, self.step_state.data_ptr(), stream))
            else:
                capi.check(L.as_adam_step(*args, self.t, stream))


    @torch.no_grad()
    def step_range(self, grad_flat, e_begin, e_end, advance):
        """The update of the flat elements [e_begin, e_end) only (e_begin a multiple of 4): used when the gradient
        arrives in chunks (dist.allreduce_and_step).  All chunks of one step see the same step count; `advance` goes
        with the last one."""
        if self.step_state is None:
            raise RuntimeError("FusedAdam.step_range needs device_step=True")
        L = capi.lib()
        p = self.params
        dev = p["means3D"].device
        if advance:
            self.t += 1
        with torch.cuda.device(dev):
            capi.check(L.as_adam_step_range(
                int(p["means3D"].shape[0]), 1 if p["log_scales"].shape[1] == 1 else 0, p["means3D"].data_ptr(),
                p["rgb_colors"].data_ptr(), p["unnorm_rotations"].data_ptr(), p["logit_opacities"].data_ptr(),
                p["log_scales"].data_ptr(), grad_flat.data_ptr(), self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(),
                self.lrs, self.betas[0], self.betas[1], self.eps, self.step_state.data_ptr(), int(e_begin), int(e_end),
                int(bool(advance)), torch.cuda.current_stream(dev).cuda_stream))

    @property
    def n_flat(self):
        """Flat elements in use: (11 + 1 or 3) per Gaussian."""
        p = self.params
        return (12 if p["log_scales"].shape[1] == 1 else 14) * int(p["means3D"].shape[0])


class FusedPoseAdam:
    """torch.optim.Adam over cam_unnorm_rots / cam_trans as the tracking loop configures it
    (initialize_optimizer(tracking=True): default betas and eps 1e-8; lrs 1e-3 / 4e-3,
    config/splatam/online_habitat_sim.py:37-45), one small kernel with the step count on the device, so that a
    pose-only iteration (get_loss(tracking=True) + backward + step) never touches the host and can be replayed as a
    CUDA graph.  `step()` reads the `.grad` of the two tensors, as torch's optimizers do."""

    def __init__(self, params, lrs, betas=(0.9, 0.999), eps=1e-8):
        self.params = params
        self.lr_rots, self.lr_trans = float(lrs["cam_unnorm_rots"]), float(lrs["cam_trans"])
        self.betas, self.eps = betas, eps
        r, t = params["cam_unnorm_rots"], params["cam_trans"]
        assert r.is_contiguous() and t.is_contiguous() and r.dtype == t.dtype == torch.float32
        n = r.numel() + t.numel()
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=r.device)
        self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=r.device)
        self.step_state = torch.zeros(8, dtype=torch.float32, device=r.device)

    @torch.no_grad()
    def step(self):
        r, t = self.params["cam_unnorm_rots"], self.params["cam_trans"]
        gr, gt = r.grad, t.grad
        if gr is None or gt is None:
            raise RuntimeError("FusedPoseAdam.step(): the camera tensors have no gradient (run backward first)")
        gr, gt = gr.contiguous(), gt.contiguous()
        with torch.cuda.device(r.device):
            capi.check(capi.lib().as_pose_adam_step(
                r.data_ptr(), gr.data_ptr(), r.numel(), self.lr_rots, t.data_ptr(), gt.data_ptr(), t.numel(),
                self.lr_trans, self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(), self.betas[0], self.betas[1],
                self.eps, self.step_state.data_ptr(), torch.cuda.current_stream(r.device).cuda_stream))


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
