"""Shared test helpers: record decoding and the oracle run over a sequence."""
import numpy as np

import np_oracle as no
import stereo.vision_b200.synth as synth


def decode_cloud(rec_i64: np.ndarray):
    """uint64 records (as int64) -> dict(x, y, d, rgb [n,3])."""
    r = rec_i64.astype(np.uint64)
    lo = (r & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    hi = (r >> np.uint64(32)).astype(np.uint32)
    return {
        "x": (lo & 0x7FF).astype(np.int32), "y": ((lo >> 11) & 0x7FF).astype(np.int32),
        "d": ((lo >> 22) & 0xFF).astype(np.uint8),
        "rgb": np.stack([hi & 0xFF, (hi >> 8) & 0xFF, (hi >> 16) & 0xFF], axis=1).astype(np.uint8),
        "spare": ((lo >> 30) | ((hi >> 24) << 2)).astype(np.int32),      # unused record bits, must be 0
    }


def decode_mcloud(rec_i32: np.ndarray):
    lo = rec_i32.astype(np.uint32)
    return {"x": (lo & 0x7FF).astype(np.int32), "y": ((lo >> 11) & 0x7FF).astype(np.int32),
            "d": ((lo >> 22) & 0xFF).astype(np.uint8)}


def oracle_sequence(disp, bgr, carmask, seeds=None, indices=None, prev=None, **kw):
    """Run the numpy oracle frame by frame, chaining the filled disparity.  Returns list of dicts."""
    out = []
    for t in range(len(disp)):
        ind = None if indices is None else (indices[0][t], indices[1][t])
        o = no.run_frame(disp[t], bgr[t], prev, carmask, seed=None if seeds is None else seeds[t],
                         indices=ind, **kw)
        prev = o["filled"]
        out.append(o)
    return out


def check_frame(res, t, o, abc_rtol=1e-9):
    """Compare frame t of a BatchResult (already on host as dict of numpy arrays) with oracle dict o."""
    N, M, K, R = [int(v) for v in res["counts"][t]]
    assert np.array_equal(res["filled"][t], o["filled"]), "filled disparity"
    c, mc = o["cloud"], o["mcloud"]
    assert N == len(c["d"]) and M == len(mc["d"]), ("counts", N, len(c["d"]), M, len(mc["d"]))
    g = decode_cloud(res["cloud"][t][:N])
    assert np.array_equal(g["y"], c["sites"][:, 0]) and np.array_equal(g["x"], c["sites"][:, 1]), "cloud order"
    assert np.array_equal(g["d"], c["d"]), "cloud d"
    assert np.array_equal(g["rgb"], c["rgb"]), "cloud rgb"
    assert not g["spare"].any(), "spare record bits"
    gm = decode_mcloud(res["mcloud"][t][:M])
    assert np.array_equal(gm["y"], mc["sites"][:, 0]) and np.array_equal(gm["x"], mc["sites"][:, 1]), "mcloud order"
    assert np.array_equal(gm["d"], mc["d"]), "mcloud d"
    pl = res["planes"][t]
    if o["abc"] is None:
        assert pl[7] == 0.0 and int(pl[6]) == -1 and K == 0 and R == 0
        assert not res["road_mask"][t].any()
        return
    assert pl[7] == 1.0
    assert int(pl[6]) == o["best"], ("best trial", int(pl[6]), o["best"])
    # a coefficient that is a rounding residue of zero (a triple with equal disparities along x) carries no relative
    # accuracy in the reference's inv() either: the absolute floor is 1e-13 of the vector's norm
    np.testing.assert_allclose(pl[:3], o["abc"], rtol=abc_rtol, atol=1e-13 * float(np.linalg.norm(o["abc"])))
    fin = np.isfinite(o["errors"])
    assert np.array_equal(np.isfinite(res["errors"][t]), fin)
    np.testing.assert_allclose(res["errors"][t][fin], o["errors"][fin], rtol=1e-9)
    rd = o["road"]
    assert K == rd["near_n"], ("near", K, rd["near_n"])
    assert np.array_equal(res["hist"][t][:no.HUE_BINS].astype(np.int64), rd["hist"]), "hue histogram"
    assert R == rd["road_n"], ("road", R, rd["road_n"])
    assert np.array_equal(res["road_mask"][t], rd["mask"]), "road mask"


def to_host(res):
    return {k: getattr(res, k).cpu().numpy() for k in
            ("filled", "counts", "planes", "errors", "hist", "road_mask", "cloud", "mcloud")}


def sanitise_cases():
    """(name, input image, threshold mask, golden record) of tests/golden/sanitise.npz: inputs are
    stored bit-packed ({0,255} images), the small case brings its own threshold mask."""
    import os
    from stereo.vision_b200.masks import default_road_threshold_mask
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sanitise.npz")
    with np.load(path, allow_pickle=False) as z:
        g = {k: z[k] for k in z.files}
    out = []
    for i in range(int(g["n_frames"])):
        p = "f%d_" % i
        r = {k[len(p):]: v for k, v in g.items() if k.startswith(p)}
        if "in_tmask" in r:
            tm = r["in_tmask"]
        else:
            tm = default_road_threshold_mask()
        H, W = tm.shape
        img = np.unpackbits(r["in_bits"])[:H * W].reshape(H, W) * np.uint8(255)
        out.append((str(r["name"]), img, tm, r))
    return out


def preprocess_cases():
    """(name, left BGR, right BGR, golden record) of tests/golden/preprocess.npz, inputs regenerated from seeds
    exactly as oracle/make_golden.py:preprocess_inputs builds them, plus the fixture's global tables."""
    import os
    import stereo.vision_b200.synth as synth
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "preprocess.npz")
    with np.load(path, allow_pickle=False) as z:
        g = {k: z[k] for k in z.files}
    ins = []
    for t in range(2):
        _, bgr = synth.make_frame(0, t)
        ins.append((bgr, np.roll(bgr, -12, axis=1)))
    rng = np.random.default_rng(4242)
    low = (100 + rng.integers(0, 9, (544, 1024, 3))).astype(np.uint8)
    ins.append((low, low[:, ::-1].copy()))
    flat = np.full((544, 1024, 3), 77, np.uint8)
    ins.append((flat, (flat // 2).astype(np.uint8)))
    small = rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
    ins.append((small, small[::-1].copy()))
    out = []
    for i, (a, b) in enumerate(ins):
        p = "f%d_" % i
        out.append((str(g[p + "name"]), a, b, {k[len(p):]: v for k, v in g.items() if k.startswith(p)}))
    return out, g


def option_cases():
    """inputs of tests/golden/options.npz rebuilt as oracle/make_golden.py:option_fixture builds them"""
    import os
    import stereo.vision_b200.synth as synth
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "options.npz")
    with np.load(path, allow_pickle=False) as z:
        g = {k: z[k] for k in z.files}
    rng = np.random.default_rng(808)
    imgs = []
    d, _ = synth.make_frame(4, 0)
    imgs.append(d)
    e = rng.integers(0, 6, (544, 1024), dtype=np.uint8)
    e[10] = 0; e[11] = 1; e[12, :500] = 0; e[13] = 255; e[14, ::2] = 0
    imgs.append(e)
    imgs.append(rng.integers(0, 256, (37, 53), dtype=np.uint8) * (rng.random((37, 53)) < 0.6))
    return [np.ascontiguousarray(i, dtype=np.uint8) for i in imgs], g


class RedZone:
    """Device buffers with 4 KB of sentinel bytes on both sides (compute-sanitizer is not available on the GPU
    pool): alloc() hands out a tensor view, check() asserts that no kernel wrote outside any of them."""
    PAD = 4096

    def __init__(self, device):
        import torch
        self.torch, self.device, self.bufs = torch, device, []

    def alloc(self, shape, dtype, fill=None):
        torch = self.torch
        n = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
        n_al = (n + 15) // 16 * 16
        raw = torch.full((self.PAD + n_al + self.PAD,), 0xA5, dtype=torch.uint8, device=self.device)
        view = raw[self.PAD:self.PAD + n].view(dtype).view(*shape)
        if fill is not None:
            view.fill_(fill)
        self.bufs.append((raw, n))
        return view

    def check(self):
        for raw, n in self.bufs:
            head, tail = raw[:self.PAD], raw[self.PAD + n:]
            assert bool((head == 0xA5).all()) and bool((tail == 0xA5).all()), "a kernel wrote outside its buffer"


def same_cycle(a, b):
    """two vertex lists describe the same oriented polygon (equal up to rotation)"""
    a, b = ([tuple(int(v) for v in p) for p in np.asarray(q).reshape(-1, 2)] for q in (a, b))
    if len(a) != len(b):
        return False
    if not a:
        return True
    if b[0] not in a:
        return False
    k = a.index(b[0])
    return a[k:] + a[:k] == b


def obstacle_cases():
    """(name, cleaned image, road points, golden record) for tests/golden/obstacles.npz: the cleaned images are
    rebuilt with the oracle from the sanitise fixture's inputs; plus the fixture's random hull cases."""
    import os
    import np_oracle as no
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "obstacles.npz")
    with np.load(path, allow_pickle=False) as z:
        g = {k: z[k] for k in z.files}
    by_name = {c[0]: c for c in sanitise_cases()}
    out = []
    for i in range(int(g["n_cases"])):
        p = "c%d_" % i
        name = str(g[p + "name"])
        _, img, tm, _ = by_name[name]
        cleaned, pts = no.sanitise_road_image(img, tm)
        out.append((name, cleaned, pts, {k[len(p):]: v for k, v in g.items() if k.startswith(p)}))
    rnd = [(g["r%d_pts" % i], str(g["r%d_fill_sha" % i]), tuple(int(v) for v in g["r%d_centre" % i])) for i in range(int(g["n_random"]))]
    return out, rnd


def sgbm_inputs():
    """(name, grayL, grayR, crop_disparity) of tests/golden/sgbm.npz: synthetic pairs of several shapes plus inputs
    that stress ties and saturation (flat images, independent noise, a pair with a single row)."""
    rng = np.random.default_rng(4242)
    cases = []
    for name, seq, t, shape in [("full_a", 0, 0, (544, 1024)), ("full_b", 5, 9, (544, 1024)), ("mid", 1, 2, (120, 400)),
                                ("narrow", 2, 1, (64, 150)), ("short", 3, 3, (9, 300)), ("odd", 4, 4, (77, 333))]:
        big = shape if shape[0] >= 120 else (120, 400)           # small cases are crops of a 120x400 pair
        L, R = synth.make_stereo_pair(seq, t, big)
        L, R = (np.ascontiguousarray(a[big[0] - shape[0]:, :shape[1]]) for a in (L, R))
        cases.append((name, L, R, False))
    L, R = synth.make_stereo_pair(6, 6, (544, 1024))
    cases.append(("full_crop", L, R, True))
    cases.append(("flat", np.full((40, 200), 90, np.uint8), np.full((40, 200), 90, np.uint8), False))
    cases.append(("noise", rng.integers(0, 256, (48, 260), dtype=np.uint8), rng.integers(0, 256, (48, 260), dtype=np.uint8), False))
    hi = (rng.integers(0, 2, (60, 280), dtype=np.uint8) * 255)
    cases.append(("binary", hi, np.roll(hi, -7, axis=1).copy(), False))
    L, R = synth.make_stereo_pair(7, 0, (120, 400))
    cases.append(("one_row", np.ascontiguousarray(L[100:101]), np.ascontiguousarray(R[100:101]), False))
    return cases


def sgbm_golden():
    import os
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sgbm.npz")
    with np.load(path, allow_pickle=False) as z:
        return {k: z[k] for k in z.files}
