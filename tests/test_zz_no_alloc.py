"""svb_compact_cloud_static: the entry point that allocates nothing (its scratch is a file-scope device array of the
library, no counterpart in the reference).  Called through the C-ABI and
held, bit for bit (integer work: records and counts), to svb_compact_cloud on the same inputs -- two calls in a row, the
second on a smaller input so that the array's stale contents would show.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _inputs(torch, dev, n, seed):
    """n cloud records (x | y << 11 | d << 22, R | G << 8 | B << 16 in the high word), distances, a hue histogram"""
    rng = np.random.default_rng(seed)
    lo = rng.integers(0, 1024, n).astype(np.uint64) | (rng.integers(0, 544, n).astype(np.uint64) << np.uint64(11)) \
        | (rng.integers(1, 255, n).astype(np.uint64) << np.uint64(22))
    hi = rng.integers(0, 1 << 24, n).astype(np.uint64)
    cloud = torch.from_numpy((lo | (hi << np.uint64(32))).view(np.int64)).to(dev)
    dist = torch.from_numpy(rng.random(n) * 0.1).to(dev)
    hist = torch.from_numpy(rng.integers(0, 25, 1024).astype(np.int32)).to(dev)
    return cloud, dist, hist


def _both(torch, N, lib, dev, cloud, dist, hist, n, mode, thr):
    """(records, count) from svb_compact_cloud with caller scratch and from the no-allocation entry point"""
    st = N.stream_ptr()
    nb = int(lib.svb_compact_scratch_bytes(n))
    scratch = torch.empty((max(nb, 16),), dtype=torch.uint8, device=dev)
    out_a = torch.full((max(n, 1),), -1, dtype=torch.int64, device=dev)
    out_b = torch.full((max(n, 1),), -1, dtype=torch.int64, device=dev)
    cnt_a = torch.full((1,), -7, dtype=torch.int32, device=dev)
    cnt_b = torch.full((1,), -7, dtype=torch.int32, device=dev)
    N.check(lib.svb_compact_cloud(N.ptr(cloud), n, mode, N.ptr(dist), N.ptr(hist), thr, N.ptr(out_a), N.ptr(cnt_a),
                                  N.ptr(scratch), scratch.numel(), st), "svb_compact_cloud")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(dev)[0]
    N.check(lib.svb_compact_cloud_static(N.ptr(cloud), n, mode, N.ptr(dist), N.ptr(hist), thr, N.ptr(out_b), N.ptr(cnt_b), st),
            "svb_compact_cloud_static")
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info(dev)[0]
    return (out_a.cpu().numpy(), int(cnt_a.item())), (out_b.cpu().numpy(), int(cnt_b.item())), free0 - free1


def test_static_scratch_compaction_equals_caller_scratch(cuda):
    import torch
    import stereo.vision_b200._native as N
    lib = N.load()
    with torch.cuda.device(cuda):
        grew = []
        # two calls in a row per mode: 300 001 records (147 block counts), then 5 000 (3 block counts, the rest stale)
        for mode, thr in ((0, 0.05), (1, 10.0)):
            for n, seed in ((300001, 1), (5000, 2)):
                cloud, dist, hist = _inputs(torch, cuda, n, seed)
                (ra, ca), (rb, cb), d = _both(torch, N, lib, cuda, cloud, dist, hist, n, mode, thr)
                grew.append(d)
                assert 0 < ca < n and ca == cb, (mode, n, ca, cb)
                assert np.array_equal(ra, rb), (mode, n)                      # kept records in order, untouched tail (-1)
                # ... and both are the stable compaction itself
                c = cloud.cpu().numpy()
                if mode == 0:
                    keep = dist.cpu().numpy() < thr
                else:
                    import np_oracle as no
                    hi = (c.view(np.uint64) >> np.uint64(32)).astype(np.uint32)
                    bins = no.hue_bin(hi & 0xFF, (hi >> 8) & 0xFF, (hi >> 16) & 0xFF)
                    keep = hist.cpu().numpy()[bins] > 10
                assert cb == int(keep.sum()) and np.array_equal(rb[:cb], c[keep])
        assert all(g == 0 for g in grew[1:]), "device memory in use grew across a call that must not allocate: %s" % grew
        # empty input, and the limit
        cloud, dist, hist = _inputs(torch, cuda, 16, 3)
        cnt = torch.full((1,), -7, dtype=torch.int32, device=cuda)
        out = torch.empty((16,), dtype=torch.int64, device=cuda)
        assert lib.svb_compact_cloud_static(N.ptr(cloud), 0, 0, N.ptr(dist), N.ptr(hist), 0.05, N.ptr(out), N.ptr(cnt), N.stream_ptr()) == 0
        assert int(cnt.item()) == 0
        too_many = N.COMPACT_STATIC_MAX_N + 1                    # rejected before anything is launched or dereferenced
        rc = lib.svb_compact_cloud_static(N.ptr(cloud), too_many, 0, N.ptr(dist), N.ptr(hist), 0.05, N.ptr(out), N.ptr(cnt), N.stream_ptr())
        assert rc == N.SVB_E_LIMIT and b"maximum" in lib.svb_error_string(rc)
        assert lib.svb_compact_cloud_static(N.ptr(cloud), 16, 2, N.ptr(dist), N.ptr(hist), 0.05, N.ptr(out), N.ptr(cnt), N.stream_ptr()) == N.SVB_E_ARG
        torch.cuda.synchronize()


def test_static_scratch_at_its_maximum(cuda):
    """n = SVB_COMPACT_STATIC_MAX_N exactly: every entry of the device array is used"""
    import torch
    import stereo.vision_b200._native as N
    lib = N.load()
    with torch.cuda.device(cuda):
        n = N.COMPACT_STATIC_MAX_N
        cloud, dist, hist = _inputs(torch, cuda, n, 4)
        (ra, ca), (rb, cb), _ = _both(torch, N, lib, cuda, cloud, dist, hist, n, 0, 0.03)
        assert ca == cb and 0 < cb < n and np.array_equal(ra, rb)
        assert cb == int((dist.cpu().numpy() < 0.03).sum())
