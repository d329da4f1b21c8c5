"""`_C`: the three extension entry points of the reference, re-implemented over the C ABI.

Same names, positional order and return layout as the pybind module of the reference
(dgr/ext.cpp:15-18; signatures dgr/rasterize_points.h:18-66; shim dgr/rasterize_points.cu:35-221),
so the reference's diff_gaussian_rasterization/__init__.py logic drives it unchanged.  This module
only allocates torch tensors and forwards raw device pointers + the current CUDA stream to
libactivesplat_b200.so; there is no computation and no fallback here.

Differences a caller can observe:
  * colors may be [P,3] (reference NUM_CHANNELS) or [P,6] (fused rgb | z,1,z^2 bundle);
  * kernels run on torch's current stream instead of the legacy default stream;
  * the three scratch tensors have a different (private) layout -- they are opaque bytes to Python.
"""
import ctypes
import threading

import torch

from . import capi

# Note on non-contiguous inputs: the reference's backward reads `scales` and `rotations` through a bare data_ptr()
# (dgr/rasterize_points.cu:173,175) although its forward used .contiguous() (:102,:104).  SplaTAM's anisotropic path
# passes rotations whose memory is column-major (F.normalize of quat_mult's torch.stack([...]).T, slam_helpers.py:28,135),
# so the stock reference differentiates w.r.t. scrambled quaternions there.  This shim always hands the kernels the
# values the forward used (what the math says, what the oracle does, and what the reference does for contiguous
# inputs); see DESIGN.md section 2.1 and tests/test_parity_gpu.py::test_noncontiguous_rotations_*.


def _dev_ptr(t, name, keep):
    """Device pointer of an optional tensor; empty tensors mean 'not provided' (rasterizer __init__.py:197-207)."""
    if t is None or t.numel() == 0:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if t.dtype != torch.float32 or not t.is_contiguous():
        t = t.contiguous().float()
    keep.append(t)
    return t.data_ptr()


class _Scratch:
    """Allocator callbacks: the library asks for bytes, torch's caching allocator provides them.

    One instance per device, created once: building ctypes callbacks on every call is slow, and a per-call object
    whose closures point back at it is a reference cycle -- its scratch tensors (tens of MB) would then live until the
    cyclic collector runs, and the caching allocator would keep growing in the meantime."""

    _per_device = {}

    def __init__(self, device):
        self.device = device
        self.tensors = {}
        self.error = None
        self.lock = threading.Lock()  # one as_forward at a time per device through this shim
        tensors = self.tensors
        state = self

        def make(tag):
            def alloc(_ctx, nbytes):
                try:
                    t = torch.empty(int(nbytes), dtype=torch.uint8, device=device)
                    tensors[tag] = t
                    return t.data_ptr()
                except Exception as e:  # surfaced after the C call returns AS_ERR_ALLOC
                    state.error = e
                    return None
            return capi.AllocFn(alloc)

        self.geom, self.binning, self.img = make("geom"), make("binning"), make("img")

    @classmethod
    def for_device(cls, device):
        key = (device.type, device.index)
        s = cls._per_device.get(key)
        if s is None:
            s = cls._per_device[key] = cls(device)
        return s

    def take(self, tag):
        """Hand the tensor to the caller (nothing stays referenced here)."""
        t = self.tensors.pop(tag, None)
        return t if t is not None else torch.empty(0, dtype=torch.uint8, device=self.device)


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                        prefiltered, debug):
    """RasterizeGaussiansCUDA (rasterize_points.cu:35-119) ->
    (num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, depth, opacity)."""
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    L = capi.lib()
    dev = means3D.device
    P, H, W = int(means3D.size(0)), int(image_height), int(image_width)
    C = int(colors.size(1)) if colors.dim() == 2 and colors.numel() > 0 else 3
    keep = []
    with torch.cuda.device(dev):
        f32 = dict(dtype=torch.float32, device=dev)
        out_color = torch.empty((C, H, W), **f32)
        out_depth = torch.empty((1, H, W), **f32)
        out_opacity = torch.empty((1, H, W), **f32)
        radii = torch.empty((P,), dtype=torch.int32, device=dev)
        scratch = _Scratch.for_device(dev)
        a = capi.AsForwardArgs()
        a.P, a.D, a.M, a.channels = P, int(degree), int(sh.size(1)) if sh.dim() > 1 and sh.size(0) != 0 else 0, C
        a.background = _dev_ptr(background, "background", keep)
        a.width, a.height = W, H
        a.means3D = _dev_ptr(means3D, "means3D", keep)
        a.shs = None
        a.colors_precomp = _dev_ptr(colors, "colors_precomp", keep)
        a.opacities = _dev_ptr(opacity, "opacities", keep)
        a.scales = _dev_ptr(scales, "scales", keep)
        a.scale_modifier = float(scale_modifier)
        a.rotations = _dev_ptr(rotations, "rotations", keep)
        a.cov3D_precomp = _dev_ptr(cov3D_precomp, "cov3D_precomp", keep)
        a.viewmatrix = _dev_ptr(viewmatrix, "viewmatrix", keep)
        a.projmatrix = _dev_ptr(projmatrix, "projmatrix", keep)
        a.cam_pos = _dev_ptr(campos, "campos", keep)
        a.tan_fovx, a.tan_fovy = float(tan_fovx), float(tan_fovy)
        a.prefiltered = int(bool(prefiltered))
        a.out_color, a.out_depth, a.out_opacity = out_color.data_ptr(), out_depth.data_ptr(), out_opacity.data_ptr()
        a.radii = radii.data_ptr() if P > 0 else None
        a.debug = int(bool(debug))
        stream = torch.cuda.current_stream(dev).cuda_stream
        with scratch.lock:
            scratch.tensors.clear()
            scratch.error = None
            rc = L.as_forward(ctypes.byref(a), scratch.geom, scratch.binning, scratch.img, None, stream)
            bufs = (scratch.take("geom"), scratch.take("binning"), scratch.take("img"))
            err = scratch.error
        if rc == capi.AS_ERR_ALLOC and err is not None:
            raise err
        capi.check(rc)
    return (rc, out_color, radii, bufs[0], bufs[1], bufs[2], out_depth, out_opacity)


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                 cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh,
                                 degree, campos, geomBuffer, R, binningBuffer, imageBuffer, debug):
    """RasterizeGaussiansBackwardCUDA (rasterize_points.cu:121-200) ->
    (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations)."""
    L = capi.lib()
    dev = means3D.device
    P = int(means3D.size(0))
    H, W = int(dL_dout_color.size(1)), int(dL_dout_color.size(2))
    C = int(dL_dout_color.size(0))
    M = int(sh.size(1)) if sh.dim() > 1 and sh.size(0) != 0 else 0
    keep = []
    with torch.cuda.device(dev):
        f32 = dict(dtype=torch.float32, device=dev)
        has_sr = scales.numel() > 0 and rotations.numel() > 0
        # every row is written by the kernel (zeros for culled Gaussians): no fill kernels needed
        dL_dmeans3D = torch.empty((P, 3), **f32)
        dL_dmeans2D = torch.empty((P, 3), **f32)
        dL_dcolors = torch.empty((P, C), **f32)
        dL_dopacity = torch.empty((P, 1), **f32)
        dL_dcov3D = torch.empty((P, 6), **f32)
        dL_dsh = torch.zeros((P, M, 3), **f32)
        dL_dscales = torch.empty((P, 3), **f32) if has_sr else torch.zeros((P, 3), **f32)
        dL_drotations = torch.empty((P, 4), **f32) if has_sr else torch.zeros((P, 4), **f32)
        if P != 0:
            a = capi.AsBackwardArgs()
            a.P, a.D, a.M, a.R, a.channels = P, int(degree), M, int(R), C
            a.background = _dev_ptr(background, "background", keep)
            a.width, a.height = W, H
            a.means3D = _dev_ptr(means3D, "means3D", keep)
            a.shs = None
            a.colors_precomp = _dev_ptr(colors, "colors_precomp", keep)
            a.scales = _dev_ptr(scales, "scales", keep)
            a.scale_modifier = float(scale_modifier)
            a.rotations = _dev_ptr(rotations, "rotations", keep)
            a.cov3D_precomp = _dev_ptr(cov3D_precomp, "cov3D_precomp", keep)
            a.viewmatrix = _dev_ptr(viewmatrix, "viewmatrix", keep)
            a.projmatrix = _dev_ptr(projmatrix, "projmatrix", keep)
            a.cam_pos = _dev_ptr(campos, "campos", keep)
            a.tan_fovx, a.tan_fovy = float(tan_fovx), float(tan_fovy)
            a.radii = radii.data_ptr() if radii is not None and radii.numel() > 0 else None
            a.geom_buffer = geomBuffer.data_ptr() if geomBuffer.numel() else None
            a.binning_buffer = binningBuffer.data_ptr() if binningBuffer.numel() else None
            a.image_buffer = imageBuffer.data_ptr() if imageBuffer.numel() else None
            a.dL_dpix = _dev_ptr(dL_dout_color, "dL_dout_color", keep)
            a.dL_dmean2D = dL_dmeans2D.data_ptr()
            a.dL_dconic = None
            a.dL_dopacity = dL_dopacity.data_ptr()
            a.dL_dcolor = dL_dcolors.data_ptr()
            a.dL_dmean3D = dL_dmeans3D.data_ptr()
            a.dL_dcov3D = dL_dcov3D.data_ptr()
            a.dL_dsh = None
            a.dL_dscale = dL_dscales.data_ptr() if has_sr else None
            a.dL_drot = dL_drotations.data_ptr() if has_sr else None
            a.debug = int(bool(debug))
            stream = torch.cuda.current_stream(dev).cuda_stream
            capi.check(L.as_backward(ctypes.byref(a), stream))
    return (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations)


def mark_visible(means3D, viewmatrix, projmatrix):
    """markVisible (rasterize_points.cu:202-221) -> bool[P]."""
    L = capi.lib()
    dev = means3D.device
    P = int(means3D.size(0))
    keep = []
    with torch.cuda.device(dev):
        present = torch.zeros((P,), dtype=torch.bool, device=dev)
        if P != 0:
            capi.check(L.as_mark_visible(P, _dev_ptr(means3D, "means3D", keep), _dev_ptr(viewmatrix, "viewmatrix", keep),
                                         _dev_ptr(projmatrix, "projmatrix", keep), present.data_ptr(),
                                         torch.cuda.current_stream(dev).cuda_stream))
    return present
