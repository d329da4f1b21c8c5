"""Writes tests/golden/ref_params_small.npz with the REFERENCE's own save_params
(src/mapper/splatam/utils/common_utils.py:37-44) on a seeded 7-Gaussian map, so the loader in
activesplat_b200/checkpoint.py is pinned on a file the reference produced.  Runs only where /root/reference exists
(this container).  The reference module imports cv2 and imgviz at the top for other helpers; neither is installed
here and save_params uses neither, so empty stand-in modules satisfy the import."""
import os
import sys
import tempfile
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
for name in ("cv2", "imgviz"):
    if name not in sys.modules:
        m = types.ModuleType(name)
        m.depth2rgb = None
        sys.modules[name] = m
sys.path.insert(0, "/root/reference/src/mapper/splatam/utils")
import common_utils as ref  # noqa: E402


def make_map(seed=3, P=7, T=4):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    params = {"means3D": r(P, 3), "rgb_colors": torch.rand(P, 3, generator=g), "unnorm_rotations": r(P, 4),
              "logit_opacities": r(P, 1), "log_scales": r(P, 1) - 4.0, "cam_unnorm_rots": r(1, 4, T),
              "cam_trans": r(1, 3, T)}
    # what SplaTAM.post_processing adds (src/mapper/splatam/__init__.py:553-570)
    params["timestep"] = torch.arange(P).float() % T
    params["intrinsics"] = np.array([[320.0, 0, 319.5], [0, 320.0, 239.5], [0, 0, 1]], np.float32)
    params["w2c"] = np.eye(4, dtype=np.float32)
    params["org_width"] = 640
    params["org_height"] = 480
    params["gt_w2c_all_frames"] = np.stack([np.eye(4, dtype=np.float32)] * T, 0)
    params["keyframe_time_indices"] = np.array([0, 2])
    return params


if __name__ == "__main__":
    with tempfile.TemporaryDirectory() as d:
        ref.save_params(make_map(), d)
        data = open(os.path.join(d, "params.npz"), "rb").read()
    open(os.path.join(HERE, "ref_params_small.npz"), "wb").write(data)
    print("wrote", os.path.join(HERE, "ref_params_small.npz"), len(data), "bytes")
