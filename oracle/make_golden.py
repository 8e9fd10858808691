"""TEST INFRASTRUCTURE -- generates tests/golden/*.npz by running the UNMODIFIED reference
(/root/reference/functions.py, imported by oracle/ref_harness.py) on seeded synthetic inputs.

Run in the build container only (the reference tree does not exist on the GPU box):

    python oracle/make_golden.py            # all fixtures, ~4 min on 8 cores

The fixtures are the only pin the oracle has (the reference ships no tests).  Inputs are
not stored: they are regenerated from (sequence id, frame id) by stereo.vision_b200/synth.py, whose
output is itself pinned here by a hash.  Large outputs are stored as sha256 digests plus compact
forms (bit-packed road mask, histogram bins, counts, plane).
"""
import hashlib
import multiprocessing as mp
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import ref_harness as rh  # noqa: E402
import stereo.vision_b200.synth as synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def sha(a) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def frame_record(o, d, bgr):
    """Compact, comparable form of one ref_harness.run_frame result."""
    rec = {
        "in_disp_sha": sha(d), "in_bgr_sha": sha(bgr),
        "filled_sha": sha(o["filled"]), "masked_disp_sha": sha(o["masked_disp"]),
        "N": len(o["points_xyz"]), "M": len(o["mask_xyz"]),
        "xyz_sha": sha(o["points_xyz"]), "rgb_sha": sha(o["points_rgb"]), "mask_xyz_sha": sha(o["mask_xyz"]),
        "xyz_head": o["points_xyz"][:64].copy(), "xyz_tail": o["points_xyz"][-64:].copy(),
        "has_plane": o["abc"] is not None,
        "abc": o["abc"] if o["abc"] is not None else np.full(3, np.nan),
        "best_error": o["best_error"], "trials_logged": o["trial_log_n"],
        "near_n": o["near_n"], "road_n": o["road_n"],
        "hist_bins": rh.hist_to_bins(o["hist"]),
        "pts_sha": sha(o["plane_pts_int"]), "pts_head": o["plane_pts_int"][:64].copy(),
        "road_mask_bits": np.packbits(o["road_mask"] != 0), "road_mask_sha": sha(o["road_mask"]),
        "road_mask_values": np.unique(o["road_mask"]),
    }
    if "diffs" in o:
        rec["diffs_head"] = o["diffs"][:256].copy()
        rec["diffs_sum"] = float(np.sum(o["diffs"]))
    return rec


def save(name, frames, **meta):
    flat = dict(meta)
    flat["n_frames"] = len(frames)
    for i, r in enumerate(frames):
        for k, v in r.items():
            flat["f%d_%s" % (i, k)] = v
    path = os.path.join(OUT, name)
    np.savez_compressed(path, **flat)
    print("wrote %s (%.1f KB)" % (name, os.path.getsize(path) / 1024))


def sequence_fixture(name, seq, n_frames, step=2, n_samples=600, trials=600):
    f = rh.load_reference(step=step, n_samples=n_samples)
    frames, prev = [], None
    for t in range(n_frames):
        d, bgr = synth.make_frame(seq, t)
        t0 = time.perf_counter()
        o = rh.run_frame(f, d, bgr, prev, synth.ransac_seed(seq, t), trials=trials)
        prev = o["filled"]
        r = frame_record(o, d, bgr)
        r["ref_seconds"] = time.perf_counter() - t0
        frames.append(r)
        print("  %s frame %d: N=%d M=%d near=%d road=%d  %.1fs" % (name, t, r["N"], r["M"], r["near_n"], r["road_n"], r["ref_seconds"]))
    save(name, frames, seq=seq, step=step, n_samples=n_samples, trials=trials, shape=np.array([544, 1024]))


def small_fixture():
    """A 64x96 frame with its own masks (module globals swapped, functions.py:32,35) and full arrays."""
    f = rh.load_reference()
    rng = np.random.default_rng(77)
    H, W = 64, 96
    d = rng.integers(0, 40, size=(H, W), dtype=np.uint8)
    d[rng.random((H, W)) < 0.25] = 0
    prev = rng.integers(0, 40, size=(H, W), dtype=np.uint8)
    bgr = rng.integers(0, 256, size=(H, W, 3), dtype=np.uint8)
    bgr[:, :, 1] = bgr[:, :, 0] // 2 + 60                      # a few populated hue bins
    cm = (rng.random((H, W)) < 0.8).astype(np.uint8) * 255
    rh.install_masks(f, carmask=cm, black=np.zeros((H, W), np.uint8))
    # the sample size is the literal 600 (functions.py:286); a 32x48 lattice has at most 1536 sites
    o = rh.run_frame(f, d, bgr, prev, seed=5, trials=50, point_threshold=0.5, road_color_thresh=2)
    r = frame_record(o, d, bgr)
    r.update(in_disp=d, in_prev=prev, in_bgr=bgr, in_carmask=cm, filled=o["filled"], xyz=o["points_xyz"],
             rgb=o["points_rgb"], mask_xyz=o["mask_xyz"], pts=o["plane_pts_int"], road_mask=o["road_mask"],
             diffs=o.get("diffs", np.zeros(0)))
    save("small_64x96.npz", [r], seed=5, trials=50, n_samples=600, point_threshold=0.5, colour_threshold=2)


def noplane_fixture():
    """Fewer masked points than the sample size: RANSAC -> (None, None), black mask."""
    f = rh.load_reference()
    d, bgr = synth.make_frame(2, 0)
    cm = np.zeros((544, 1024), np.uint8)
    cm[200:210, 100:160] = 255
    rh.install_masks(f, carmask=cm)
    o = rh.run_frame(f, d, bgr, None, seed=1)
    r = frame_record(o, d, bgr)
    save("noplane.npz", [r], seq=2)


def _hue_chunk(r):
    f = rh.load_reference()
    out = np.empty((256, 256), dtype=np.uint16)
    rr = np.uint8(r)
    for g in range(256):
        gg = np.uint8(g)
        for b in range(256):
            out[g, b] = int(round(float(f.BGRtoHSVHue((rr, gg, np.uint8(b)))) * 1000.0))
    return r, out


def hue_fixture():
    """Reference hue key of ALL 2^24 colours (functions.py:73-78 with np.uint8 channels), as bins."""
    table = np.empty((256, 256, 256), dtype=np.uint16)
    with mp.get_context("fork").Pool(os.cpu_count()) as pool:
        for r, t in pool.imap_unordered(_hue_chunk, range(256)):
            table[r] = t
    # keep the hash of the full table and, verbatim, every 17th red plane (1 MB raw, compresses well)
    np.savez_compressed(os.path.join(OUT, "hue_bins.npz"), sha=sha(table), planes=np.arange(0, 256, 17),
                        table_planes=table[::17].copy(), populated=len(np.unique(table)))
    print("wrote hue_bins.npz, %d distinct bins, sha %s" % (len(np.unique(table)), sha(table)[:16]))


def reproject_fixture():
    """x' for every (x, d) and y' for every (y, d) through the reference's own projection and
    reprojection helpers (functions.py:178-209) and the caller's int32 cast (stereovision.py:112);
    the step-1 variant of the loader gives odd coordinates too, the unmodified one is cross-checked."""
    out = {}
    for step in (1, 2):
        f = rh.load_reference(step=step)
        xt = np.full((1024, 256), -1, dtype=np.int32)
        yt = np.full((544, 256), -1, dtype=np.int32)
        for d in range(1, 256):
            row = np.full((2, 1025), d, dtype=np.uint8)           # y = 0 only, x = 0..1023
            pts = f.projectDisparityTo3d(row, 128)
            pp = np.array(f.project3DPointsTo2DImagePoints(pts), np.int32)
            xt[0:1024:step, d] = pp[:, 0]
            col = np.full((545, 2), d, dtype=np.uint8)            # x = 0 only, y = 0..543
            pts = f.projectDisparityTo3d(col, 128)
            pp = np.array(f.project3DPointsTo2DImagePoints(pts), np.int32)
            yt[0:544:step, d] = pp[:, 1]
        out[step] = (xt, yt)
    x1, y1 = out[1]
    x2, y2 = out[2]
    assert np.array_equal(x1[::2], x2[::2]) and np.array_equal(y1[::2], y2[::2])
    dx = (np.arange(1024)[:, None] - x1)[:, 1:]
    dy = (np.arange(544)[:, None] - y1)[:, 1:]
    assert set(np.unique(dx)) <= {0, 1} and set(np.unique(dy)) <= {0, 1}
    np.savez_compressed(os.path.join(OUT, "reproject_table.npz"), dx_bits=np.packbits(dx.astype(np.uint8)),
                        dy_bits=np.packbits(dy.astype(np.uint8)), moved_x=int(dx.sum()), moved_y=int(dy.sum()))
    print("wrote reproject_table.npz: %d (x,d) and %d (y,d) pairs land one pixel before" % (dx.sum(), dy.sum()))


def sanitise_inputs():
    """Seeded road images for the cleaning fixture: the three reference road masks of seq0 (read back
    from their own fixture), speckle at several densities over a dense patch, and shapes built to sit
    on the edges of ParticleCleansing's 70 px^2 rule (rings around 7..11 px holes, blobs at the border)."""
    out = []
    z = np.load(os.path.join(OUT, "seq0_step2_s600.npz"))
    for i in range(int(z["n_frames"])):
        out.append(("seq0_f%d" % i, np.unpackbits(z["f%d_road_mask_bits" % i])[:544 * 1024].reshape(544, 1024) * np.uint8(255)))
    rng = np.random.default_rng(2024)
    for dens in (0.002, 0.03, 0.3):
        img = ((rng.random((544, 1024)) < dens) * 255).astype(np.uint8)
        img[330:520, 150:900] |= ((rng.random((190, 750)) < 0.55) * 255).astype(np.uint8)
        out.append(("speckle_%g" % dens, img))
    img = np.zeros((544, 1024), np.uint8)
    for j, hole in enumerate((7, 8, 9, 10, 11, 13)):
        y0, x0 = 300, 60 + 150 * j
        img[y0:y0 + 60, x0:x0 + 60] = 255
        img[y0 + 25:y0 + 25 + hole, x0 + 25:x0 + 25 + hole] = 0          # a hole of hole x hole px
    img[400:430, 0:30] = 255; img[514:544, 500:530] = 255; img[350:380, 994:1024] = 255   # touching borders
    img[450:452, 300:302] = 255; img[460:469, 340:349] = 255; img[480:505, 400:425] = 255   # small blobs
    out.append(("shapes", img))
    return out


def sanitise_fixture():
    f = rh.load_reference()
    recs = []
    for name, img in sanitise_inputs():
        t0 = time.perf_counter()
        cleaned, pts = rh.run_sanitise(f, img)
        recs.append({"name": name, "in_bits": np.packbits(img != 0), "in_sha": sha(img),
                     "out_bits": np.packbits(cleaned != 0), "out_sha": sha(cleaned), "out_values": np.unique(cleaned),
                     "n_pts": len(pts), "pts_sha": sha(pts), "pts_head": pts[:64].copy(), "pts_tail": pts[-64:].copy(),
                     "ref_seconds": time.perf_counter() - t0})
        print("  sanitise %-14s in=%d out=%d  %.2fs" % (name, int((img != 0).sum()), len(pts), recs[-1]["ref_seconds"]))
    # a small frame with its own threshold mask (module global swapped, functions.py:33)
    rng = np.random.default_rng(9)
    img = ((rng.random((70, 101)) < 0.35) * 255).astype(np.uint8)
    img[20:60, 10:90] |= ((rng.random((40, 80)) < 0.7) * 255).astype(np.uint8)
    tm = np.zeros((70, 101), np.uint8); tm[15:, :] = 128; tm[40:, 5:97] = 255
    cleaned, pts = rh.run_sanitise(f, img, tm)
    recs.append({"name": "small_70x101", "in_bits": np.packbits(img != 0), "in_sha": sha(img), "in_tmask": tm,
                 "out_bits": np.packbits(cleaned != 0), "out_sha": sha(cleaned), "out_values": np.unique(cleaned),
                 "n_pts": len(pts), "pts_sha": sha(pts), "pts_head": pts[:64].copy(), "pts_tail": pts[-64:].copy(),
                 "ref_seconds": 0.0, "pts": pts})
    save("sanitise.npz", recs)


def preprocess_inputs():
    """Seeded colour pairs for the front-end fixture: two synthetic frames (the right image is the left
    shifted by 12 px), a low-contrast pair (few grey levels: the equalisation LUT is mostly steps), a
    single-colour image (equalizeHist's degenerate branch) and a 37x53 pair."""
    out = []
    for t in range(2):
        _, bgr = synth.make_frame(0, t)
        out.append(("seq0_f%d" % t, bgr, np.roll(bgr, -12, axis=1)))
    rng = np.random.default_rng(4242)
    low = (100 + rng.integers(0, 9, (544, 1024, 3))).astype(np.uint8)
    out.append(("low_contrast", low, low[:, ::-1].copy()))
    flat = np.full((544, 1024, 3), 77, np.uint8)
    out.append(("flat", flat, (flat // 2).astype(np.uint8)))
    small = rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
    out.append(("small_37x53", small, small[::-1].copy()))
    return out


def preprocess_fixture():
    f = rh.load_reference()
    flat = {}
    # gamma tables straight from the reference's expression (functions.py:62-65) through cv2.LUT on a ramp
    ramp = np.arange(256, dtype=np.uint8).reshape(16, 16)
    gammas = np.array([1.4, 1.0, 2.2, 0.5, 1.8])
    flat["gammas"] = gammas
    flat["gamma_tables"] = np.stack([f.gammaChange(ramp, float(g)).reshape(256) for g in gammas])
    # BGR2GRAY over all 2^24 colours, through the reference's greyscale() minus equalisation: cv2 itself
    import cv2
    r, g, b = np.meshgrid(np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8), indexing="ij")
    cube = np.stack([b, g, r], axis=-1).reshape(4096, 4096, 3)          # index = r*65536 + g*256 + b, channels B,G,R
    gray = cv2.cvtColor(cube, cv2.COLOR_BGR2GRAY)
    flat["gray_all_sha"] = sha(gray)
    flat["gray_r128"] = gray.reshape(256, 256, 256)[128].copy()          # the red = 128 plane verbatim
    recs = []
    for name, imgL, imgR in preprocess_inputs():
        gl, gr, yl, yr = rh.run_preprocess(f, imgL, imgR)
        recs.append({"name": name, "inL_sha": sha(imgL), "inR_sha": sha(imgR), "gammaL_sha": sha(gl), "gammaR_sha": sha(gr),
                     "grayL_sha": sha(yl), "grayR_sha": sha(yr),
                     "grayL_hist": np.bincount(yl.ravel(), minlength=256), "grayR_hist": np.bincount(yr.ravel(), minlength=256),
                     "grayL_head": yl[:2, :64].copy()})
        print("  preprocess %-13s %s" % (name, sha(yl)[:12]))
    small = preprocess_inputs()[-1]
    gl, gr, yl, yr = rh.run_preprocess(f, small[1], small[2])
    recs[-1].update(inL=small[1], inR=small[2], gammaL=gl, grayL=yl, grayR=yr)
    flat["n_frames"] = len(recs)
    for i, r_ in enumerate(recs):
        for k, v in r_.items():
            flat["f%d_%s" % (i, k)] = v
    path = os.path.join(OUT, "preprocess.npz")
    np.savez_compressed(path, **flat)
    print("wrote preprocess.npz (%.1f KB)" % (os.path.getsize(path) / 1024))


def option_fixture():
    """fillAltDisparity on seeded disparities (rows without any value, rows of ones, a ragged
    width) and getNormalVectorLine on a grid of base points x planes x disparities."""
    f = rh.load_reference()
    flat = {}
    rng = np.random.default_rng(808)
    imgs = []
    d, _ = synth.make_frame(4, 0)
    imgs.append(d)
    e = rng.integers(0, 6, (544, 1024), dtype=np.uint8)            # many values below 2, small means
    e[10] = 0; e[11] = 1; e[12, :500] = 0; e[13] = 255; e[14, ::2] = 0
    imgs.append(e)
    imgs.append(rng.integers(0, 256, (37, 53), dtype=np.uint8) * (rng.random((37, 53)) < 0.6))
    flat["n_fill"] = len(imgs)
    for i, im in enumerate(imgs):
        im = np.ascontiguousarray(im, dtype=np.uint8)
        out = rh.run_fill_alt(f, im)
        if im.size <= 4096:
            flat["fill%d_in" % i] = im
        flat["fill%d_in_sha" % i] = sha(im)
        flat["fill%d_shape" % i] = np.array(im.shape)
        flat["fill%d_out_sha" % i] = sha(out)
        flat["fill%d_out_rows" % i] = out[:16].copy()
        print("  fillAlt %d: %s" % (i, sha(out)[:12]))
    # the normal glyph
    planes = [np.array([0.002, 0.9, 0.03]), np.array([-0.05, 1.1, 0.01]), np.array([0.0, 0.0, 0.5]), np.array([0.3, -0.2, 0.04])]
    pts, res = [], []
    disp = np.zeros((544, 1024), np.uint8)
    for k, abc in enumerate(planes):
        for (x, y) in [(0, 0), (512, 300), (1023, 543), (474, 262), (100, 500), (900, 400)]:
            for dv in (0, 1, 7, 40, 128, 255):
                disp[y, x] = dv
                r = rh.run_normal_line(f, (x, y), abc, disp)
                pts.append((k, x, y, dv))
                res.append(r if isinstance(r, tuple) else (-(2 ** 31), {"ValueError": 1, "OverflowError": 2}.get(r, 3)))
    flat["normal_planes"] = np.stack(planes)
    flat["normal_cases"] = np.array(pts, dtype=np.int64)
    flat["normal_results"] = np.array(res, dtype=np.int64)
    path = os.path.join(OUT, "options.npz")
    np.savez_compressed(path, **flat)
    print("wrote options.npz (%.1f KB), %d normal-line cases, %d raise" % (os.path.getsize(path) / 1024, len(res),
                                                                            sum(1 for r in res if r[0] == -(2 ** 31))))


def obstacle_fixture():
    """hull of the road points, hull mask, obstacle mask and hull centre on the cleaned images of
    the sanitise fixture (stereovision.py:170-200), plus the hull-mask rule against cv2 on random hulls."""
    import cv2
    import np_oracle as no
    f = rh.load_reference()
    flat, n = {}, 0
    for name, img in sanitise_inputs():
        cleaned, pts = rh.run_sanitise(f, img)
        if len(pts) == 0:
            continue
        hull, hmask, obst, centre = rh.run_obstacles(f, cleaned, pts)
        flat["c%d_name" % n] = name
        flat["c%d_hull" % n] = hull.astype(np.int32)
        flat["c%d_hull_mask_sha" % n] = sha(hmask)
        flat["c%d_obstacle_sha" % n] = sha(obst)
        flat["c%d_obstacle_px" % n] = int((obst != 0).sum())
        flat["c%d_centre" % n] = np.array(centre)
        print("  obstacles %-14s hull %d vertices, %d obstacle px, centre %s" % (name, len(hull), int((obst != 0).sum()), centre))
        n += 1
    flat["n_cases"] = n
    # the span rule of np_oracle.hull_fill against drawContours on random hulls, segments and points
    rng = np.random.default_rng(77)
    cases, ok = [], 0
    for trial in range(300):
        k = [30, 3, 2, 1][trial % 4]
        pts = np.stack([rng.integers(0, 1024, k), rng.integers(0, 544, k)], 1).astype(np.int32)
        cvh = cv2.convexHull(pts)
        ref = cv2.drawContours(np.zeros((544, 1024), np.uint8), [cvh], 0, 255, -100)
        ok += int(np.array_equal(ref, no.hull_fill(no.convex_hull(pts), (544, 1024))))
        if trial < 24:
            cases.append((pts, sha(ref), tuple(int(v) for v in f.getCenterPoint(cvh))))
    assert ok == 300, ok
    flat["n_random"] = len(cases)
    for i, (pts, h, c) in enumerate(cases):
        flat["r%d_pts" % i], flat["r%d_fill_sha" % i], flat["r%d_centre" % i] = pts, h, np.array(c)
    path = os.path.join(OUT, "obstacles.npz")
    np.savez_compressed(path, **flat)
    print("wrote obstacles.npz (%.1f KB)" % (os.path.getsize(path) / 1024))


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    which = sys.argv[1:] or ["seq", "s400", "step1", "small", "noplane", "reproject", "hue", "sanitise", "preprocess", "options", "obstacles"]
    if "seq" in which:
        sequence_fixture("seq0_step2_s600.npz", 0, 3)
    if "s400" in which:
        sequence_fixture("seq6_step2_s400.npz", 6, 2, n_samples=400)
    if "step1" in which:
        sequence_fixture("seq7_step1_s600.npz", 7, 1, step=1)
    if "small" in which:
        small_fixture()
    if "noplane" in which:
        noplane_fixture()
    if "reproject" in which:
        reproject_fixture()
    if "hue" in which:
        hue_fixture()
    if "sanitise" in which:
        sanitise_fixture()
    if "preprocess" in which:
        preprocess_fixture()
    if "options" in which:
        option_fixture()
    if "obstacles" in which:
        obstacle_fixture()
