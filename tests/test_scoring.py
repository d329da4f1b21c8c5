"""Clustering after the GetOpacity renders: get_invisibility_clusters (src/mapper/__init__.py:92-117).

CPU: the oracle restatement against outputs of the reference's own function (tests/golden/ref_clusters.npz).
GPU: as_dbscan_grid through activesplat_b200.scoring against sklearn's labels (bit-exact) and the same golden file."""
import os

import numpy as np
import pytest
import torch

from oracle import scoring_oracle as so

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_clusters.npz")


def _golden():
    with np.load(GOLDEN) as z:
        return {k: z[k] for k in z.files}


def test_oracle_matches_the_reference_function_outputs():
    g = _golden()
    for s in g["seeds"]:
        centers, sums = so.get_invisibility_clusters(so.blob_image(int(s)), 30)
        assert np.array_equal(np.array(centers, dtype=np.float64).reshape(-1, 2), g[f"centers_{s}"])
        assert np.array_equal(np.array(sums, dtype=np.float32), g[f"sums_{s}"])


def test_oracle_labels_image_conventions():
    img = np.zeros((20, 30), np.float32)
    lab = so.dbscan_labels_image(img, 0.3, 5, 10)
    assert (lab == -2).all()                      # no samples at all (the reference returns [], [])
    img[5:9, 5:9] = 1.0                           # 16 samples, every one within eps of >= 10 others
    img[15, 25] = 1.0                             # an isolated sample: noise
    lab = so.dbscan_labels_image(img, 0.3, 5, 10)
    assert (lab[5:9, 5:9] == 0).all() and lab[15, 25] == -1 and (lab == -2).sum() == 20 * 30 - 17


@pytest.mark.gpu
def test_dbscan_grid_labels_are_sklearns():
    from activesplat_b200 import scoring
    imgs = [so.blob_image(s) for s in range(8)]
    imgs.append(np.zeros((75, 180), np.float32))                       # empty
    imgs.append(np.ones((75, 180), np.float32))                        # one cluster filling the image
    chk = np.indices((75, 180)).sum(0) % 2                             # checkerboard: dense but every pixel a sample edge
    imgs.append(chk.astype(np.float32))
    two = np.zeros((75, 180), np.float32)                              # two clusters bridged by a thin line of border pixels
    two[10:20, 10:20] = 1.0
    two[10:20, 27:37] = 1.0
    two[15, 20:27] = 1.0
    imgs.append(two)
    batch = torch.from_numpy(np.stack(imgs)).cuda()
    for thr, eps, ms in ((0.3, 5.0, 10), (0.8, 5.0, 25), (0.3, 3.0, 4)):
        labels, n, si, sv = scoring.dbscan_grid(batch, thr, eps=eps, min_samples=ms, max_clusters=128)
        labels, n, si, sv = labels.cpu().numpy(), n.cpu().numpy(), si.cpu().numpy(), sv.cpu().numpy()
        for b, img in enumerate(imgs):
            want = so.dbscan_labels_image(img, thr, eps, ms)
            assert np.array_equal(labels[b], want), (b, thr, eps, ms)          # bit-exact, noise and border pixels included
            k = int(want.max()) + 1 if (want >= 0).any() else 0
            assert n[b] == k
            for c in range(k):
                rows, cols = np.where(want == c)
                assert si[b, c, 0] == len(rows) and si[b, c, 1] == rows.sum() and si[b, c, 2] == cols.sum()
                assert abs(sv[b, c] - img[rows, cols].astype(np.float64).sum()) <= 1e-9 * max(1.0, len(rows))


@pytest.mark.gpu
def test_get_invisibility_clusters_matches_reference_outputs():
    from activesplat_b200 import scoring
    g = _golden()
    for s in g["seeds"]:
        img = torch.from_numpy(so.blob_image(int(s))).cuda()
        centers, sums = scoring.get_invisibility_clusters(img, 30)
        want_c, want_s = g[f"centers_{s}"], g[f"sums_{s}"]
        assert len(sums) == len(want_s)
        assert np.allclose(np.array(centers).reshape(-1, 2), want_c, rtol=0, atol=1e-12)   # exact integer sums / counts
        assert np.allclose(np.array(sums, dtype=np.float64), want_s, rtol=1e-5)            # float32 pairwise sum vs double


@pytest.mark.gpu
def test_local_invisibility_path_from_a_panorama():
    """src/mapper/splatam/__init__.py:793-832 after the renders, against the same steps done with numpy + the oracle."""
    from activesplat_b200 import scoring
    rng = np.random.default_rng(11)
    inv = np.zeros((150, 360), np.float32)
    inv[40:90, 250:330] = rng.uniform(0.5, 1.0, (50, 80)).astype(np.float32)   # a big unseen region off-centre
    inv[100:110, 20:40] = 0.6
    opacity = torch.from_numpy(1.0 - inv).cuda()
    c2w = np.eye(4)
    total, pose = scoring.local_invisibility_from_panorama(opacity, c2w, 30)
    assert abs(total - float((1.0 - (1.0 - inv)).sum())) <= 1e-3 * total
    small = 0.25 * (inv[0::2, 0::2] + inv[0::2, 1::2] + inv[1::2, 0::2] + inv[1::2, 1::2])   # INTER_AREA, factor 0.5
    centers, sums = so.get_invisibility_clusters(small, 30)
    c = centers[int(np.argmax(sums))]
    h, v = np.deg2rad(c[1] / 0.5 - 60), np.deg2rad(c[0] / 0.5 - 75)
    want = scoring.rot_axis(scoring.rot_axis(c2w, 'y', h), 'x', v)
    assert pose is not None and np.allclose(pose, want, atol=1e-9)
    # nothing unseen: no pose
    total0, pose0 = scoring.local_invisibility_from_panorama(torch.ones(150, 360).cuda(), c2w, 30)
    assert total0 == 0.0 and pose0 is None


@pytest.mark.gpu
def test_get_local_invisibility_end_to_end_against_the_oracle_pipeline():
    """LOCAL GetOpacity branch (src/mapper/splatam/__init__.py:762-838): three look-around renders + panorama + clusters
    + best pose on the device, against the same pipeline built from the CPU oracles (oracle rasterizer for the
    renders, sklearn for the clusters)."""
    from activesplat_b200 import scenes, scoring, splatam
    from oracle import splatam_oracle as spo
    dev = torch.device("cuda:0")
    # a room with one wall missing: a large unseen region on one side
    params_cpu, _ = scenes.s_room(6000, 1, seed=41, scale_mult=6.0)
    keep = params_cpu["means3D"][:, 0] < 3.0
    params_cpu = {k: v[keep].contiguous() for k, v in params_cpu.items()}
    params = {k: v.to(dev) for k, v in params_cpu.items()}
    c2w = np.eye(4)
    c2w[:3, 3] = [0.3, 0.0, -0.4]
    total, pose = scoring.get_local_invisibility(params, c2w, 30)
    # oracle pipeline on the CPU
    ops = []
    for v in scoring.look_around_views(c2w):
        cam = splatam.setup_camera(v.width, v.height, v.k(), v.w2c, device="cpu")
        rv, drv = spo.get_rendervars(params_cpu, torch.as_tensor(v.w2c).float())
        _, _, opacity, _ = spo.render(cam, splatam.with_background(cam, [1.0, 1.0, 1.0]), rv, drv)
        ops.append(opacity[0].numpy())
    inv = 1.0 - np.hstack(ops)
    want_total = float(inv.sum())
    assert abs(total - want_total) <= 2e-4 * want_total
    assert want_total > 100
    small = 0.25 * (inv[0::2, 0::2] + inv[0::2, 1::2] + inv[1::2, 0::2] + inv[1::2, 1::2])
    centers, sums = so.get_invisibility_clusters(small, 30)
    assert len(sums) > 0
    c = centers[int(np.argmax(sums))]
    h, v_ = np.deg2rad(c[1] / 0.5 - 60), np.deg2rad(c[0] / 0.5 - 75)
    if abs(h) > np.deg2rad(15) or abs(v_) > np.deg2rad(15):
        want = scoring.rot_axis(scoring.rot_axis(c2w, 'y', h), 'x', v_)
        assert pose is not None and np.allclose(pose, want, atol=2e-3)   # cluster centre moves by < 0.1 px
    else:
        assert pose is None


def test_matrix_to_pose_is_the_w_positive_unit_quaternion():
    """src/utils/gui_utils.py:138-152 (quaternion.from_rotation_matrix), checked against scipy's conversion."""
    from scipy.spatial.transform import Rotation
    from activesplat_b200.getopacity import matrix_to_pose
    rng = np.random.default_rng(0)
    for _ in range(20):
        R = Rotation.random(random_state=int(rng.integers(1 << 30)))
        M = np.eye(4)
        M[:3, :3], M[:3, 3] = R.as_matrix(), rng.normal(size=3)
        p = matrix_to_pose(M)
        q = R.as_quat()
        q = q if q[3] >= 0 else -q
        assert np.allclose(p.orientation_xyzw, q, atol=1e-12) and np.array_equal(p.position, M[:3, 3])


@pytest.mark.gpu
def test_get_opacity_local_and_global_renders():
    from activesplat_b200 import getopacity as go, scenes, scoring
    dev = torch.device("cuda:0")
    params_cpu, _ = scenes.s_room(6000, 1, seed=41, scale_mult=6.0)
    keep = params_cpu["means3D"][:, 0] < 3.0
    params = {k: v[keep].contiguous().to(dev) for k, v in params_cpu.items()}
    c2w = np.eye(4)
    c2w[:3, 3] = [0.3, 0.0, -0.4]
    resp = go.handle_get_opacity(go.GetOpacityRequest(arrived_flag=False), params, c2w)
    total, pose = scoring.get_local_invisibility(params, c2w, 30)
    assert resp.targets_frustums_volume == [0.0] and len(resp.targets_frustums) == 1
    assert abs(resp.targets_frustums_invisibility[0] - total) <= 1e-3 * total
    if pose is None:
        assert np.array_equal(resp.targets_frustums[0].orientation_xyzw, np.zeros(4))
    else:
        want = go.matrix_to_pose(pose @ go.OPENCV_TO_OPENGL)
        assert np.allclose(resp.targets_frustums[0].orientation_xyzw, want.orientation_xyzw, atol=1e-6)
        assert np.allclose(resp.targets_frustums[0].position, c2w[:3, 3])


# ---- get_convexhull_volume (GLOBAL branch, src/mapper/__init__.py:8-90) -------------------------------------------
HULL_GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_hull.npz")


def test_hull_oracle_matches_the_reference_function_outputs():
    """oracle/hull_oracle.py (own dilation, border following, incremental hull) vs outputs of the reference's
    get_convexhull_volume (OpenCV + scikit-learn + SciPy) on seeded panoramas (tests/golden/make_hull_golden.py)."""
    from oracle import hull_oracle as ho
    with np.load(HULL_GOLDEN) as z:
        g = {k: z[k] for k in z.files}
    for s in g["seeds"]:
        depth, inv = ho.panorama(int(s))
        li, lv = ho.get_convexhull_volume(depth, inv)
        assert abs(lv - g[f"volume_{s}"]) <= 1e-9 * max(abs(g[f"volume_{s}"]), 1e-12), (s, lv, g[f"volume_{s}"])
        # the reference sums invisibility * volume in float32
        assert abs(li - g[f"invisibility_{s}"]) <= 1e-5 * max(abs(g[f"invisibility_{s}"]), 1e-12), (s, li)


def test_hull_oracle_pieces():
    """Element shape, single pixels, degenerate point sets."""
    from oracle import hull_oracle as ho
    assert ho.ellipse_rows() == [0, 4, 5, 6, 6, 7, 7, 7, 7, 7, 6, 6, 5, 4, 0]
    m = np.zeros((40, 50), bool)
    m[20, 25] = True
    d = ho.dilate(m)
    assert d.sum() == sum(2 * dx + 1 for dx in ho.ellipse_rows()) and d[13, 25] and d[20, 18] and not d[13, 24]
    assert ho.outer_contour_simple(m) == [(25, 20)]
    m[0, 0] = True                                                      # a component cut by the image corner
    assert ho.dilate(m)[:8, :8].sum() > 30
    cube = np.array([[x, y, z] for x in (0, 1) for y in (0, 1) for z in (0, 1)], float)
    assert abs(ho.convex_hull_volume(cube) - 1.0) < 1e-12
    assert abs(ho.convex_hull_volume(np.vstack([cube, [[0.5, 0.5, 0.5]], [[0.2, 0.9, 0.1]]])) - 1.0) < 1e-12
    assert ho.convex_hull_volume(np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [1, 1, 0], [0.3, 0.3, 0]], float)) == 0.0


@pytest.mark.gpu
def test_cluster_hull_volumes_match_oracle_and_reference():
    """as_dbscan_grid + as_cluster_hull_volumes (scoring.get_convexhull_volume) on a batch of panoramas: per panorama
    the same (invisibility, volume) as the oracle and as the reference's own function (golden file)."""
    from activesplat_b200 import scoring
    from oracle import hull_oracle as ho
    with np.load(HULL_GOLDEN) as z:
        g = {k: z[k] for k in z.files}
    seeds = [int(s) for s in g["seeds"]] + [100, 101, 102]
    pans = [ho.panorama(s) for s in seeds]
    # a panorama whose clusters all lie on "no surface" pixels, one without clusters, one that is a single cluster
    d0, i0 = ho.panorama(200)
    d0[:] = ho.EMPTY_DEPTH
    pans.append((d0, i0))
    pans.append((np.full((150, 360, 1), 2.0, np.float32), np.zeros((150, 360), np.float32)))
    pans.append((ho.panorama(201)[0], np.ones((150, 360), np.float32)))
    depth = torch.from_numpy(np.stack([p[0][..., 0] for p in pans])).cuda()
    inv = torch.from_numpy(np.stack([p[1] for p in pans])).cuda()
    li, lv, n_cl, n_pts = scoring.get_convexhull_volume(depth, inv)
    li, lv = li.cpu().numpy(), lv.cpu().numpy()
    assert int(n_pts.min()) >= 0                                        # no contour overflowed the vertex buffer
    for b, (d, i) in enumerate(pans):
        want_i, want_v = ho.get_convexhull_volume(d, i)
        assert abs(lv[b] - want_v) <= 1e-9 * max(abs(want_v), 1e-9), (b, lv[b], want_v)
        assert abs(li[b] - want_i) <= 1e-5 * max(abs(want_i), 1e-9), (b, li[b], want_i)
    for k, s in enumerate(g["seeds"]):
        assert abs(lv[k] - g[f"volume_{s}"]) <= 1e-9 * max(abs(g[f"volume_{s}"]), 1e-12)
        assert abs(li[k] - g[f"invisibility_{s}"]) <= 1e-5 * max(abs(g[f"invisibility_{s}"]), 1e-12)
    assert lv[len(seeds)] == 0.0 and li[len(seeds)] == 0.0 and lv[len(seeds) + 1] == 0.0
    # one panorama alone gives the same numbers as inside the batch
    li1, lv1, _, _ = scoring.get_convexhull_volume(depth[3], inv[3])
    assert abs(float(lv1[0]) - lv[3]) <= 1e-12 * max(lv[3], 1e-12)


@pytest.mark.gpu
def test_get_opacity_global_handler_matches_oracle_pipeline():
    """GLOBAL GetOpacity (visualizer.py:2183-2195 + :991-1001): all nodes in one pass on the device vs, per node, the
    oracle pipeline on the same renders (panorama of the three look-around views -> hull_oracle.get_convexhull_volume)."""
    from activesplat_b200 import dist as asd, getopacity as go, scenes, scoring
    from oracle import hull_oracle as ho
    dev = torch.device("cuda:0")
    params_cpu, _ = scenes.s_room(8000, 1, seed=41, scale_mult=6.0)
    keep = (params_cpu["means3D"][:, 0] < 2.5) | (params_cpu["means3D"][:, 2] > 3.0)   # a map with unexplored sides
    params = {k: v[keep].contiguous().to(dev) for k, v in params_cpu.items()}
    c2w = np.eye(4)
    c2w[:3, 3] = [0.3, 0.0, -0.4]
    nodes = np.array([[1.0, 0.0, 0.5], [0.0, 0.0, 0.0], [-1.0, 0.0, -0.5], [2.5, 0.0, 2.0], [-3.0, 0.0, 1.0]])
    resp = go.handle_get_opacity(go.GetOpacityRequest(True, nodes, list(range(len(nodes)))), params, c2w)
    assert resp.targets_frustums == [] and len(resp.targets_frustums_invisibility) == len(nodes)
    assert resp.targets_frustums_invisibility[1] == 0.0 and resp.targets_frustums_volume[1] == 0.0   # node at the origin
    some = 0
    for n_i, pos in enumerate(nodes):
        if (pos == 0).all():
            continue
        views = scoring.look_around_views(c2w, position=pos)
        _, ops, deps = asd.score_candidate_views(params, views, n_streams=3, return_opacity=True, return_depth=True)
        inv_np = 1.0 - np.hstack([o.cpu().numpy() for o in ops])
        dep_np = np.hstack([d.cpu().numpy() for d in deps])[..., None]
        want_i, want_v = ho.get_convexhull_volume(dep_np, inv_np)
        got_i, got_v = resp.targets_frustums_invisibility[n_i], resp.targets_frustums_volume[n_i]
        assert abs(got_v - want_v) <= 1e-7 * max(abs(want_v), 1e-9), (n_i, got_v, want_v)
        assert abs(got_i - want_i) <= 1e-5 * max(abs(want_i), 1e-9), (n_i, got_i, want_i)
        some += want_v > 0
    assert some >= 2   # the scene leaves something to find
