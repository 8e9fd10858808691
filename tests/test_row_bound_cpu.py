"""near_row_ranges (pipeline.py): the lattice rows the end-to-end streamer uploads the colour image for must contain every
near-plane point the reference would find (functions.py:300-323 on the cloud of functions.py:178-198), for planes a RANSAC
run would rarely choose as well as for road planes.  CPU only: the oracle's float64 distances decide what is near."""
import numpy as np

import np_oracle as no
import stereo.vision_b200.synth as synth
from stereo.vision_b200.camera import DEFAULT_CAMERA
from stereo.vision_b200.pipeline import near_row_ranges

H, W, STEP, THR = 544, 1024, 2, 0.05


def _planes(rng):
    pl = [synth.true_plane(s, t) for s in range(4) for t in range(6)]                  # the bench's road planes
    for _ in range(120):                                                               # perturbed road planes
        pl.append(np.array([0.02, 1.65, 0.03]) * (1.0 + 0.2 * rng.normal(size=3)))
    for _ in range(150):                                                               # any direction, nine decades of |abc|
        v = rng.normal(size=3)
        pl.append(v / np.linalg.norm(v) * 10.0 ** rng.uniform(-3, 6))
    for _ in range(60):                                                                # steep / frontal / ceiling planes
        pl.append(np.array([rng.normal() * 5, rng.normal() * 0.05, rng.normal() * 0.05]) if rng.random() < 0.5 else
                  np.array([rng.normal() * 0.01, -abs(rng.normal()) * 2.0, 10.0 ** rng.uniform(-2, 1)]))
    return np.array(pl)


def test_rows_with_near_points_lie_inside_the_uploaded_range():
    rng = np.random.default_rng(7)
    cam = no.cam_of(None)
    # every disparity value occurs on every lattice row, so a row holds a near point whenever the plane allows one
    d = rng.integers(1, 255, size=(H, W), dtype=np.uint8)
    d[:, 0:2 * 254:2] = np.arange(1, 255, dtype=np.uint8)[None, :]
    cloud = no.project(d, None, STEP, None)
    ly = cloud["sites"][:, 0] // STEP
    planes = _planes(rng)
    rows8 = np.zeros((len(planes), 8))
    for i, abc in enumerate(planes):
        nrm = float(np.sqrt(abc[0] * abc[0] + abc[1] * abc[1] + abc[2] * abc[2]))
        rows8[i] = [abc[0], abc[1], abc[2], nrm, THR * nrm, 0.0, 0.0, 1.0]
    first, count = near_row_ranges(rows8, DEFAULT_CAMERA, H, W, STEP)
    held = tight = 0
    for i, abc in enumerate(planes):
        near = no.plane_distance(cloud["xyz"], abc) < THR
        if near.any():
            r = ly[near]
            assert count[i] > 0 and first[i] <= r.min() and r.max() < first[i] + count[i], (i, abc, r.min(), r.max(), first[i], count[i])
            held += 1
            tight += int(count[i] <= (r.max() - r.min() + 1) + 8)
    assert held > 150 and tight > held // 2, (held, tight)          # the bound is used, and mostly within a few rows of exact
    # invalid planes upload nothing; the bench's planes skip the rows above the horizon
    bad = rows8[:3].copy()
    bad[0, 7] = 0.0
    bad[1, 0] = np.nan
    bad[2, 4] = np.inf
    assert (near_row_ranges(bad, DEFAULT_CAMERA, H, W, STEP)[1] == 0).all()
    assert (count[:24] < 160).all() and (first[:24] > 100).all()
