"""The on-disk map format of ActiveSplat / SplaTAM: `params.npz` (+ `params<t>.npz` checkpoints).

Mirrors src/mapper/splatam/utils/common_utils.py:26-68 (`params2cpu`, `save_params`, `save_params_ckpt`) and the
keys SplaTAM.post_processing adds before saving (src/mapper/splatam/__init__.py:553-570), and adds the loader the
reference leaves to offline tools, so maps written by a real ActiveSplat run can be fed to the kernels here
(and maps optimised here can be read back by the reference's viewers).

File contents (all numpy arrays, `np.savez`, uncompressed):
  the seven optimised tensors   means3D [P,3], rgb_colors [P,3], unnorm_rotations [P,4], logit_opacities [P,1],
                                log_scales [P,1] or [P,3], cam_unnorm_rots [1,4,T], cam_trans [1,3,T]
  bookkeeping                   timestep [P], intrinsics [3,3], w2c [4,4] (first frame), org_width, org_height,
                                gt_w2c_all_frames [T,4,4], keyframe_time_indices [K]
"""
import os

import numpy as np
import torch

GAUSSIAN_KEYS = ("means3D", "rgb_colors", "unnorm_rotations", "logit_opacities", "log_scales")
CAMERA_KEYS = ("cam_unnorm_rots", "cam_trans")
META_KEYS = ("timestep", "intrinsics", "w2c", "org_width", "org_height", "gt_w2c_all_frames", "keyframe_time_indices")


def params2cpu(params):
    """common_utils.py:26-34: tensors -> contiguous CPU numpy arrays, everything else passes through."""
    res = {}
    for k, v in params.items():
        if isinstance(v, torch.Tensor):
            res[k] = v.detach().cpu().contiguous().numpy()
        else:
            res[k] = v
    return res


def save_params(output_params, output_dir):
    """common_utils.py:37-44: writes <output_dir>/params.npz."""
    to_save = params2cpu(output_params)
    os.makedirs(output_dir, exist_ok=True)
    save_path = os.path.join(output_dir, "params.npz")
    np.savez(save_path, **to_save)
    return save_path


def save_params_ckpt(output_params, output_dir, time_idx):
    """common_utils.py:61-68: writes <output_dir>/params<time_idx>.npz."""
    to_save = params2cpu(output_params)
    os.makedirs(output_dir, exist_ok=True)
    save_path = os.path.join(output_dir, "params" + str(time_idx) + ".npz")
    np.savez(save_path, **to_save)
    return save_path


def add_camera_metadata(params, variables, intrinsics, first_frame_w2c, org_width, org_height, gt_w2c_all_frames,
                        keyframe_time_indices):
    """The keys SplaTAM.post_processing puts into `params` before save_params (src/mapper/splatam/__init__.py:553-570),
    including the trim of the camera tracks to the number of frames actually seen."""
    params = dict(params)
    params["timestep"] = variables["timestep"]
    params["intrinsics"] = intrinsics.detach().cpu().numpy() if isinstance(intrinsics, torch.Tensor) else np.asarray(intrinsics)
    params["w2c"] = (first_frame_w2c.detach().cpu().numpy() if isinstance(first_frame_w2c, torch.Tensor)
                     else np.asarray(first_frame_w2c))
    params["org_width"] = org_width
    params["org_height"] = org_height
    params["gt_w2c_all_frames"] = np.stack(
        [g.detach().cpu().numpy() if isinstance(g, torch.Tensor) else np.asarray(g) for g in gt_w2c_all_frames], axis=0)
    params["keyframe_time_indices"] = np.array(keyframe_time_indices)
    n = params["gt_w2c_all_frames"].shape[0]
    for key in CAMERA_KEYS:
        if params[key].shape[-1] > n:
            params[key] = params[key][..., :n]
    return params


def load_params(path, device="cpu", requires_grad=False):
    """Read a params.npz (file or the directory holding it).

    Returns (params, meta): `params` holds the seven optimised tensors as float32 torch tensors on `device`
    (contiguous; `requires_grad` as asked), `meta` the bookkeeping arrays as numpy (scalars unwrapped).
    Unknown keys are kept in `meta`.  Raises KeyError if one of the five Gaussian tensors is missing and ValueError
    if their leading dimensions disagree."""
    if os.path.isdir(path):
        path = os.path.join(path, "params.npz")
    with np.load(path, allow_pickle=False) as z:
        arrays = {k: z[k] for k in z.files}
    for k in GAUSSIAN_KEYS:
        if k not in arrays:
            raise KeyError(f"{path}: missing '{k}'")
    P = arrays["means3D"].shape[0]
    for k in GAUSSIAN_KEYS:
        if arrays[k].shape[0] != P:
            raise ValueError(f"{path}: '{k}' has {arrays[k].shape[0]} rows, means3D has {P}")
    params, meta = {}, {}
    for k, v in arrays.items():
        if k in GAUSSIAN_KEYS or k in CAMERA_KEYS:
            t = torch.from_numpy(np.ascontiguousarray(v, dtype=np.float32)).to(device)
            params[k] = t.requires_grad_(True) if requires_grad else t
        else:
            meta[k] = v.item() if v.ndim == 0 else v
    return params, meta
