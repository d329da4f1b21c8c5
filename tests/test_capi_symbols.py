"""CPU: the C-ABI library loads and exports every symbol include/activesplat_b200.h declares.
No compute call is made here (there is no GPU in the CPU tier)."""
import ctypes
import os
import re

import helpers  # noqa: F401  (sys.path)
from activesplat_b200 import capi

HEADER = os.path.join(helpers.ROOT, "include", "activesplat_b200.h")


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(as_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    L = ctypes.CDLL(capi.LIB_PATH)
    names = _declared_functions()
    assert len(names) >= 12
    for n in names:
        assert hasattr(L, n), f"{n} declared in the header but not exported"
    # and the Python binding knows all of them
    assert set(names) == set(capi.SYMBOLS), set(names) ^ set(capi.SYMBOLS)


def test_sizes_and_layout_helpers_run_without_gpu():
    L = capi.lib()
    assert L.as_geometry_bytes(1000) > 1000 * (4 + 8 + 16 + 4 + 4 + 4 + 48)
    assert L.as_image_bytes(640, 480) >= 640 * 480 * 8 + 1200 * 8
    assert L.as_binning_bytes(1000, 6) > L.as_binning_bytes(1000, 3)
    assert L.as_sort_temp_bytes(10) > 0 and L.as_scan_temp_bytes(10) > 0
    assert [L.as_higher_msb(n) for n in (1200, 256, 1024, 80)] == [11, 9, 11, 7]
    lay = capi.AsScratchLayout()
    assert L.as_scratch_layout(100, 64, 64, 500, 3, ctypes.byref(lay)) == 0
    assert lay.keys_sorted > lay.keys_unsorted
    assert L.as_scratch_layout(100, 64, 64, 500, 4, ctypes.byref(lay)) == capi.AS_ERR_INVALID_ARG
    assert L.as_error_string(capi.AS_ERR_NO_COLORS) == b"For non-RGB, provide precomputed Gaussian colors!"


def test_python_surface_matches_reference_names():
    import diff_gaussian_rasterization as d
    assert d.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug")
    for n in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible"):
        assert callable(getattr(d._C, n))
    import torch
    r = d.GaussianRasterizer(None)
    x = torch.zeros(1, 3)
    for kw in (dict(), dict(shs=x, colors_precomp=x)):
        try:
            r(x, x, x, scales=x, rotations=x, **kw)
            assert False
        except Exception as e:
            assert "excatly one of either SHs or precomputed colors" in str(e)
    try:
        r(x, x, x, colors_precomp=x, scales=x)
        assert False
    except Exception as e:
        assert "exactly one of either scale/rotation pair" in str(e)
