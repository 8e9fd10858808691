"""Pins the oracle (oracle/np_oracle.py) to outputs of the UNMODIFIED reference stored in tests/golden
(made by oracle/make_golden.py from /root/reference/functions.py).  CPU only."""
import hashlib
import os

import numpy as np
import pytest

import np_oracle as no
import stereo.vision_b200.synth as synth
from stereo.vision_b200.masks import default_carmask

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def load(name):
    with np.load(os.path.join(GOLD, name), allow_pickle=False) as z:
        return {k: z[k] for k in z.files}


def frame(g, i):
    p = "f%d_" % i
    return {k[len(p):]: v for k, v in g.items() if k.startswith(p)}


def compare(o, r):
    """oracle frame dict o vs golden record r: bit-exact sets/orders/masks, planes to 1e-9."""
    assert sha(o["filled"]) == str(r["filled_sha"])
    c, m, rd = o["cloud"], o["mcloud"], o["road"]
    assert len(c["d"]) == int(r["N"]) and len(m["d"]) == int(r["M"])
    assert sha(c["xyz"]) == str(r["xyz_sha"]), "XYZ float64 must be bit-identical to the reference's"
    assert sha(c["rgb"]) == str(r["rgb_sha"])
    assert sha(m["xyz"]) == str(r["mask_xyz_sha"])
    if not bool(r["has_plane"]):
        assert o["abc"] is None and rd["road_n"] == 0 and not rd["mask"].any()
        return
    np.testing.assert_allclose(o["abc"], r["abc"], rtol=1e-9)
    np.testing.assert_allclose(o["errors"][o["best"]], float(r["best_error"]), rtol=1e-9)
    assert rd["near_n"] == int(r["near_n"]) and rd["road_n"] == int(r["road_n"])
    assert np.array_equal(rd["hist"], r["hist_bins"])
    assert sha(rd["pts"]) == str(r["pts_sha"]), "reprojected integer points, in order"
    assert sha(rd["mask"]) == str(r["road_mask_sha"])
    assert np.array_equal(np.packbits(rd["mask"] != 0), r["road_mask_bits"])
    np.testing.assert_allclose(rd["dist"][:256], r["diffs_head"], rtol=1e-10, atol=1e-13)


@pytest.mark.parametrize("name", ["seq0_step2_s600.npz", "seq6_step2_s400.npz", "seq7_step1_s600.npz"])
def test_sequences(name):
    g = load(name)
    seq, step, S, T = int(g["seq"]), int(g["step"]), int(g["n_samples"]), int(g["trials"])
    cm = default_carmask((544, 1024), prefer_cwd=False)
    prev = None
    for t in range(int(g["n_frames"])):
        r = frame(g, t)
        d, bgr = synth.make_frame(seq, t)
        assert sha(d) == str(r["in_disp_sha"]) and sha(bgr) == str(r["in_bgr_sha"]), "synthetic input drifted"
        o = no.run_frame(d, bgr, prev, cm, seed=synth.ransac_seed(seq, t), trials=T, n_samples=S, step=step)
        prev = o["filled"]
        compare(o, r)


def test_small_frame_with_own_masks():
    g = load("small_64x96.npz")
    r = frame(g, 0)
    o = no.run_frame(r["in_disp"], r["in_bgr"], r["in_prev"], r["in_carmask"], seed=int(g["seed"]),
                     trials=int(g["trials"]), n_samples=int(g["n_samples"]),
                     threshold=float(g["point_threshold"]), colour_threshold=int(g["colour_threshold"]))
    compare(o, r)
    assert np.array_equal(o["filled"], r["filled"]) and np.array_equal(o["cloud"]["xyz"], r["xyz"])
    assert np.array_equal(o["road"]["pts"], r["pts"]) and np.array_equal(o["road"]["mask"], r["road_mask"])


def test_no_plane():
    g = load("noplane.npz")
    r = frame(g, 0)
    d, bgr = synth.make_frame(int(g["seq"]), 0)
    cm = np.zeros((544, 1024), np.uint8)
    cm[200:210, 100:160] = 255
    o = no.run_frame(d, bgr, None, cm, seed=1)
    assert not bool(r["has_plane"])
    compare(o, r)


def test_hue_bins_all_colours():
    """a10 exhaustively: oracle hue bin of all 2^24 colours == the reference's key (hash + sampled planes)."""
    g = load("hue_bins.npz")
    h = hashlib.sha256()
    gg, bb = np.meshgrid(np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8), indexing="ij")
    planes = {int(p): i for i, p in enumerate(g["planes"])}
    for r in range(256):
        t = no.hue_bin(np.full_like(gg, r), gg, bb).astype(np.uint16)
        h.update(np.ascontiguousarray(t).tobytes())
        if r in planes:
            assert np.array_equal(t, g["table_planes"][planes[r]])
    assert h.hexdigest() == str(g["sha"])


def test_reprojection_table():
    """a13 exhaustively: every (x,d), (y,d) pair lands on the site or one pixel before, as in the reference."""
    g = load("reproject_table.npz")
    c = no.cam_of(None)
    for n, centre, key, moved in ((1024, c.cw, "dx_bits", "moved_x"), (544, c.ch, "dy_bits", "moved_y")):
        coord = np.arange(n, dtype=np.float64)[:, None]
        d = np.arange(1, 256, dtype=np.float64)[None, :]
        Z = (c.f * c.B) / d
        V = ((coord - centre) * Z) / c.f
        xyz = np.stack([V.ravel(), V.ravel(), np.broadcast_to(Z, V.shape).ravel()], axis=1)
        cam = no.Cam(c.f, c.B, centre, centre)
        _, pi = no.reproject(xyz, cam)
        delta = (np.arange(n)[:, None] - pi[:, 0].reshape(n, 255)).astype(np.uint8)
        assert np.array_equal(np.packbits(delta), g[key]) and int(delta.sum()) == int(g[moved])


def test_sanitise_road_image():
    """oracle f1 (road-mask cleaning) vs the unmodified sanitiseRoadImage, ParticleCleansing included."""
    from helpers import sanitise_cases
    cases = sanitise_cases()
    assert len(cases) == 8
    for name, img, tm, r in cases:
        assert sha(img) == str(r["in_sha"]), name
        out, pts = no.sanitise_road_image(img, tm)
        assert sha(out) == str(r["out_sha"]), name
        assert set(np.unique(out)) <= {0, 255}
        assert len(pts) == int(r["n_pts"]) and sha(pts) == str(r["pts_sha"]), name
        assert np.array_equal(pts[:64], r["pts_head"]) and np.array_equal(pts[-64:], r["pts_tail"])
        if "pts" in r:
            assert np.array_equal(pts, r["pts"])


def test_preprocess_front_end():
    """oracle f2 (gamma LUT, BGR->gray, equalizeHist) vs the unmodified preProcessImages / greyscale and cv2."""
    from helpers import preprocess_cases
    cases, g = preprocess_cases()
    for k, gamma in enumerate(g["gammas"]):
        assert np.array_equal(no.gamma_table(float(gamma)), g["gamma_tables"][k]), gamma
    r, gg, b = np.meshgrid(np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8), indexing="ij")
    gray = no.bgr_to_gray(np.stack([b, gg, r], axis=-1).reshape(4096, 4096, 3))
    assert sha(gray) == str(g["gray_all_sha"]), "BGR2GRAY over all 2^24 colours"
    assert np.array_equal(gray.reshape(256, 256, 256)[128], g["gray_r128"])
    for name, imgL, imgR, rec in cases:
        assert sha(imgL) == str(rec["inL_sha"]) and sha(imgR) == str(rec["inR_sha"]), name
        gl, gr, yl, yr = no.preprocess_pair(imgL, imgR, 1.4)
        assert sha(gl) == str(rec["gammaL_sha"]) and sha(gr) == str(rec["gammaR_sha"]), name
        assert sha(yl) == str(rec["grayL_sha"]) and sha(yr) == str(rec["grayR_sha"]), name


def test_fill_alt_and_normal_line():
    """oracle f3 vs the unmodified fillAltDisparity ('mean' option) and getNormalVectorLine"""
    from helpers import option_cases
    imgs, g = option_cases()
    assert len(imgs) == int(g["n_fill"])
    for i, im in enumerate(imgs):
        assert sha(im) == str(g["fill%d_in_sha" % i]), i
        out = no.fill_alt_disparity(im)
        assert sha(out) == str(g["fill%d_out_sha" % i]), i
        assert np.array_equal(out[:16], g["fill%d_out_rows" % i])
    planes, cases, results = g["normal_planes"], g["normal_cases"], g["normal_results"]
    raised = 0
    for (k, x, y, dv), want in zip(cases, results):
        try:
            got = no.normal_vector_line((int(x), int(y)), planes[k], int(dv))
        except (ValueError, OverflowError) as e:
            got = (-(2 ** 31), 1 if isinstance(e, ValueError) else 2)
            raised += 1
        assert tuple(int(v) for v in want) == got, (k, x, y, dv)
    assert raised == 24


def test_obstacle_stage():
    """oracle f3 (hull, hull mask, obstacle mask, hull centre) vs cv2 as stereovision.py:170-200 calls it"""
    from helpers import obstacle_cases, same_cycle
    cases, rnd = obstacle_cases()
    assert len(cases) == 7
    for name, cleaned, pts, r in cases:
        hull, hmask, obst = no.obstacle_stage(cleaned, pts)
        assert same_cycle(hull, r["hull"]), name
        assert sha(hmask) == str(r["hull_mask_sha"]) and sha(obst) == str(r["obstacle_sha"]), name
        assert no.enclosing_centre(hull) == tuple(int(v) for v in r["centre"]), name
    for pts, fill_sha, centre in rnd:
        hull = no.convex_hull(pts)
        assert sha(no.hull_fill(hull, (544, 1024))) == fill_sha
        assert no.enclosing_centre(hull) == centre


# ---------------------------------------------------------------- the matcher stage
def test_sgbm_disparity_stage():
    """oracle/sgbm_oracle.c against the unmodified functions.disparity (cv2.StereoSGBM + filterSpeckles + scaling,
    functions.py:26,104-128) on every case of tests/golden/sgbm.npz: raw int16 matcher output, speckle-filtered
    image and final uint8 disparity bit for bit."""
    import sgbm_oracle as so
    from helpers import sgbm_golden, sgbm_inputs
    g = sgbm_golden()
    cases = sgbm_inputs()
    assert [c[0] for c in cases] == [str(n) for n in g["names"]]
    for name, L, R, crop in cases:
        p = name + "_"
        assert sha(L) + sha(R) == str(g[p + "in_sha"]), "synthetic input drifted: " + name
        raw = so.sgbm_compute(L, R, 128, 21)
        assert sha(raw) == str(g[p + "raw_sha"]), name
        assert int((raw >= 0).sum()) == int(g[p + "valid"])
        assert sha(so.filter_speckles(raw.copy(), 0, 60, 40)) == str(g[p + "speckle60_sha"]), name
        spk = so.filter_speckles(raw.copy(), 0, 4000, 123)
        assert sha(spk) == str(g[p + "speckle_sha"]), name
        final = so.scale_disparity(spk, 128, crop)
        assert final.shape == tuple(g[p + "shape"]) and sha(final) == str(g[p + "final_sha"]), name
        if p + "final" in g:
            assert sha(so.disparity(L, R, 128, crop)) == str(g[p + "final_sha"])
            assert np.array_equal(final, g[p + "final"]) and np.array_equal(raw, g[p + "raw"])
