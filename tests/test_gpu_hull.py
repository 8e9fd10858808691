"""GPU: SURVEY 8(f)-3 -- the obstacle stage (stereovision.py:170-200: convexHull, hull mask, obstacle image,
getCenterPoint) through the C-ABI against cv2 as the reference calls it (tests/golden/obstacles.npz) and the oracle."""
import numpy as np
import pytest
import torch

import np_oracle as no
import stereo.vision_b200._native as N
from helpers import RedZone, obstacle_cases, same_cycle
from test_oracle_golden import sha

pytestmark = pytest.mark.gpu


def run_hull(cuda, imgs, cap=None, masks=True):
    """imgs uint8 [F,H,W] (0 / non-zero) -> hull list, centres, hull masks, obstacle masks (numpy)."""
    lib = N.load()
    imgs = np.ascontiguousarray(imgs)
    F, H, W = imgs.shape
    cap = 2 * H if cap is None else cap
    t = torch.from_numpy(imgs).to(cuda)
    rz = RedZone(cuda)
    bits = rz.alloc((F, int(lib.svb_bits_words(H, W))), torch.int32)
    N.check(lib.svb_pack_mask(N.ptr(t), F, H, W, N.ptr(bits), None, 0), "svb_pack_mask")
    sb = int(lib.svb_road_hull_scratch_bytes(F, H))
    assert sb == 4 * F * H
    scratch = rz.alloc((sb,), torch.uint8) if masks else None
    hull = rz.alloc((F, max(cap, 1), 2), torch.int32, fill=-7)
    hn = rz.alloc((F,), torch.int32)
    centre = rz.alloc((F, 2), torch.int32)
    hm = rz.alloc((F, H, W), torch.uint8, fill=3) if masks else None
    ob = rz.alloc((F, H, W), torch.uint8, fill=3) if masks else None
    N.check(lib.svb_road_hull(N.ptr(bits), F, H, W, cap, N.ptr(hull), N.ptr(hn), N.ptr(centre),
                              N.ptr(hm), N.ptr(ob), N.ptr(scratch), sb if masks else 0, 0), "svb_road_hull")
    torch.cuda.synchronize()
    rz.check()
    hn = hn.cpu().numpy()
    hull = hull.cpu().numpy()
    hulls = [hull[f, :min(hn[f], cap)] for f in range(F)]
    return hulls, hn, centre.cpu().numpy(), (hm.cpu().numpy() if masks else None), (ob.cpu().numpy() if masks else None)


def test_obstacle_stage_golden(cuda):
    cases, _ = obstacle_cases()
    by_shape = {}
    for c in cases:
        by_shape.setdefault(c[1].shape, []).append(c)
    for shape, group in by_shape.items():
        hulls, hn, centre, hm, ob = run_hull(cuda, np.stack([c[1] for c in group]))
        for i, (name, cleaned, pts, r) in enumerate(group):
            assert hn[i] == len(r["hull"]), name
            assert same_cycle(hulls[i], r["hull"]), name
            # canonical start: the lexicographically largest (x, y)
            assert tuple(hulls[i][0]) == max(map(tuple, hulls[i].tolist())), name
            assert sha(hm[i]) == str(r["hull_mask_sha"]), name
            assert sha(ob[i]) == str(r["obstacle_sha"]), name
            assert tuple(centre[i]) == tuple(int(v) for v in r["centre"]), name


def test_random_hulls_fill_and_centre(cuda):
    """the fixture's random point sets (cv2's own fill + centre), all in one batch"""
    _, rnd = obstacle_cases()
    imgs = np.zeros((len(rnd), 544, 1024), dtype=np.uint8)
    for i, (pts, _, _) in enumerate(rnd):
        imgs[i, pts[:, 1], pts[:, 0]] = 255
    hulls, hn, centre, hm, ob = run_hull(cuda, imgs)
    for i, (pts, fill_sha, c) in enumerate(rnd):
        want = no.convex_hull(pts)
        assert same_cycle(hulls[i], want), i
        assert sha(hm[i]) == fill_sha, i                     # sparse points: hull mask == fill
        assert np.array_equal(ob[i], np.where(imgs[i] != 0, 0, hm[i])), i
        assert tuple(centre[i]) == c, i


def test_shapes_and_degenerate_images(cuda):
    rng = np.random.default_rng(11)
    for H, W in ((1, 1), (1, 40), (40, 1), (7, 33), (64, 100), (130, 2048), (2048, 96)):
        imgs = np.zeros((6, H, W), dtype=np.uint8)
        imgs[1] = 255                                                        # full
        imgs[2, rng.integers(0, H), rng.integers(0, W)] = 255                # one pixel
        imgs[3, H // 2, :] = 255                                             # one row
        imgs[4, :, W // 2] = 255                                             # one column
        imgs[5] = (rng.random((H, W)) < 0.02) * 255
        k = np.arange(min(H, W))
        imgs[0, k, k] = 255                                                  # a diagonal: collinear points
        hulls, hn, centre, hm, ob = run_hull(cuda, imgs)
        for f in range(6):
            ys, xs = np.nonzero(imgs[f])
            pts = np.stack([xs, ys], 1)
            if len(pts) == 0:
                assert hn[f] == 0 and tuple(centre[f]) == (-1, -1)
                assert not hm[f].any() and not ob[f].any()
                continue
            want, whm, wob = no.obstacle_stage(imgs[f], pts)
            assert same_cycle(hulls[f], want), (H, W, f)
            assert np.array_equal(hm[f], whm) and np.array_equal(ob[f], wob), (H, W, f)
            assert tuple(centre[f]) == no.enclosing_centre(want), (H, W, f)


def test_cap_and_optional_outputs(cuda):
    cases, _ = obstacle_cases()
    name, cleaned, pts, r = cases[0]
    n = len(r["hull"])
    hulls, hn, centre, _, _ = run_hull(cuda, cleaned[None], cap=3, masks=False)
    assert hn[0] == n and len(hulls[0]) == 3
    full, _, _, _, _ = run_hull(cuda, cleaned[None], masks=False)
    assert np.array_equal(full[0][:3], hulls[0])
    lib = N.load()
    z = torch.zeros((4,), dtype=torch.int32, device=cuda)
    assert lib.svb_road_hull(N.ptr(z), 1, 1, 1, 0, None, None, None, None, None, None, 0, 0) == N.SVB_E_NULL
    assert lib.svb_road_hull(None, 1, 1, 1, 0, None, N.ptr(z), None, None, None, None, 0, 0) == N.SVB_E_NULL
    assert lib.svb_road_hull(N.ptr(z), 1, 1, 4096, 0, None, N.ptr(z), None, None, None, None, 0, 0) == N.SVB_E_SHAPE
    assert lib.svb_road_hull(N.ptr(z), 1, 1, 1, 0, N.ptr(z), N.ptr(z), None, None, None, None, 0, 0) == N.SVB_E_ARG
    m = torch.zeros((16,), dtype=torch.uint8, device=cuda)
    assert lib.svb_road_hull(N.ptr(z), 1, 1, 1, 0, None, None, None, N.ptr(m), None, None, 0, 0) == N.SVB_E_SCRATCH
    assert lib.svb_road_hull(N.ptr(z), 1, 1, 1, 0, None, None, None, N.ptr(m), None, N.ptr(z), 3, 0) == N.SVB_E_SCRATCH


def test_drop_ins_and_cleaner(cuda):
    import stereo.vision_b200.functions as f
    from stereo.vision_b200.pipeline import RoadCleaner
    from helpers import sanitise_cases
    cases, rnd = obstacle_cases()
    name, cleaned, pts, r = cases[0]
    hull, hm, ob = f.roadObstacles(cleaned)
    assert hull.shape == (len(r["hull"]), 1, 2) and hull.dtype == np.int32
    assert same_cycle(hull, r["hull"]) and sha(hm) == str(r["hull_mask_sha"]) and sha(ob) == str(r["obstacle_sha"])
    assert f.getCenterPoint(hull) == tuple(int(v) for v in r["centre"])
    assert f.getCenterPoint(rnd[0][0]) == rnd[0][2]                         # any point set, not only a hull
    with pytest.raises(ValueError):
        f.roadObstacles(np.zeros_like(cleaned))
    # batched: RoadCleaner.clean -> obstacles
    by_name = {c[0]: c for c in sanitise_cases()}
    group = [c for c in cases if c[1].shape == cleaned.shape]
    raw = np.stack([by_name[c[0]][1] for c in group])
    tm = by_name[group[0][0]][2]
    if all(np.array_equal(by_name[c[0]][2], tm) for c in group):
        rc = RoadCleaner(shape=cleaned.shape, threshold_mask=tm, max_frames=len(group), device=cuda)
        rc.clean(raw)
        out = rc.obstacles()
        torch.cuda.synchronize()
        for i, c in enumerate(group):
            n = int(out["hull_n"][i])
            assert same_cycle(out["hull"][i, :n].cpu().numpy(), c[3]["hull"]), c[0]
            assert sha(out["obstacles"][i].cpu().numpy()) == str(c[3]["obstacle_sha"]), c[0]
            assert tuple(out["centre"][i].tolist()) == tuple(int(v) for v in c[3]["centre"]), c[0]
