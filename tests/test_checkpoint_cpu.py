"""params.npz (src/mapper/splatam/utils/common_utils.py:37-68): the loader/savers in activesplat_b200/checkpoint.py
against a file written by the reference's own save_params (tests/golden/ref_params_small.npz, made by
tests/golden/make_checkpoint_golden.py)."""
import os

import numpy as np
import pytest
import torch

from activesplat_b200 import checkpoint as ck

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_params_small.npz")


def _seeded_map(seed=3, P=7, T=4):
    # same generator sequence as make_checkpoint_golden.make_map
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    return {"means3D": r(P, 3), "rgb_colors": torch.rand(P, 3, generator=g), "unnorm_rotations": r(P, 4),
            "logit_opacities": r(P, 1), "log_scales": r(P, 1) - 4.0, "cam_unnorm_rots": r(1, 4, T),
            "cam_trans": r(1, 3, T)}


def test_load_file_written_by_the_reference():
    params, meta = ck.load_params(GOLDEN)
    want = _seeded_map()
    assert set(params) == set(want)
    for k, v in want.items():
        assert params[k].dtype == torch.float32 and params[k].is_contiguous()
        assert np.array_equal(params[k].numpy(), v.numpy()), k  # bit-exact
    assert meta["org_width"] == 640 and meta["org_height"] == 480
    assert meta["w2c"].shape == (4, 4) and meta["intrinsics"].shape == (3, 3)
    assert meta["gt_w2c_all_frames"].shape == (4, 4, 4)
    assert list(meta["keyframe_time_indices"]) == [0, 2]
    assert np.array_equal(meta["timestep"], np.arange(7, dtype=np.float32) % 4)


def test_our_save_is_byte_compatible_with_the_reference_loader_side(tmp_path):
    """np.load of our file yields the same keys, dtypes, shapes and values as the reference's file."""
    with np.load(GOLDEN) as z:
        ref = {k: z[k] for k in z.files}
    mine = dict(_seeded_map())
    mine = ck.add_camera_metadata(mine, {"timestep": torch.arange(7).float() % 4},
                                  np.array([[320.0, 0, 319.5], [0, 320.0, 239.5], [0, 0, 1]], np.float32),
                                  torch.eye(4), 640, 480, [torch.eye(4)] * 4, [0, 2])
    path = ck.save_params(mine, str(tmp_path))
    assert os.path.basename(path) == "params.npz"
    with np.load(path) as z:
        got = {k: z[k] for k in z.files}
    assert set(got) == set(ref)
    for k in ref:
        assert got[k].dtype == ref[k].dtype and got[k].shape == ref[k].shape, k
        assert np.array_equal(got[k], ref[k]), k
    p2 = ck.save_params_ckpt(mine, str(tmp_path), 12)
    assert os.path.basename(p2) == "params12.npz"


def test_camera_tracks_are_trimmed_to_the_frames_seen():
    m = _seeded_map(T=9)
    out = ck.add_camera_metadata(m, {"timestep": torch.zeros(7)}, np.eye(3), np.eye(4), 64, 48, [np.eye(4)] * 5, [0])
    assert out["cam_trans"].shape[-1] == 5 and out["cam_unnorm_rots"].shape[-1] == 5


def test_loader_rejects_incomplete_or_ragged_maps(tmp_path):
    m = {k: v.numpy() for k, v in _seeded_map().items()}
    bad = dict(m)
    del bad["log_scales"]
    np.savez(tmp_path / "a.npz", **bad)
    with pytest.raises(KeyError):
        ck.load_params(str(tmp_path / "a.npz"))
    bad = dict(m)
    bad["rgb_colors"] = bad["rgb_colors"][:3]
    np.savez(tmp_path / "b.npz", **bad)
    with pytest.raises(ValueError):
        ck.load_params(str(tmp_path / "b.npz"))


def test_empty_map_round_trips(tmp_path):
    m = {"means3D": torch.zeros(0, 3), "rgb_colors": torch.zeros(0, 3), "unnorm_rotations": torch.zeros(0, 4),
         "logit_opacities": torch.zeros(0, 1), "log_scales": torch.zeros(0, 1), "cam_unnorm_rots": torch.zeros(1, 4, 1),
         "cam_trans": torch.zeros(1, 3, 1)}
    ck.save_params(m, str(tmp_path))
    params, meta = ck.load_params(str(tmp_path))
    assert params["means3D"].shape == (0, 3) and meta == {}
