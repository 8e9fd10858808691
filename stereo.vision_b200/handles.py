"""Return types of the drop-in helpers.

The reference's helpers return Python lists (a cloud is a list of [X,Y,Z,R,G,B] lists), a dict and
numpy arrays, and stereovision.py only ever takes len(), iterates, indexes or converts them with
np.array.  The handles below keep the data on the GPU as packed records and provide
exactly those operations, materialising on the host lazily and once.
"""
import ctypes as C

import numpy as np
import torch

from . import _native as N


class PointCloud:
    """Device cloud: uint64 records (x,y,d,R,G,B,hue) in the reference's append order."""

    def __init__(self, records, n, camera, has_rgb):
        self.records = records            # int64 CUDA tensor, first n entries valid
        self.n = int(n)
        self.camera = camera
        self.has_rgb = bool(has_rgb)
        self._host = None

    def __len__(self):
        return self.n

    # -- host materialisation (functions.py:191-195 values, bit-exact float64)
    def _export(self):
        if self._host is None:
            n, dev = self.n, self.records.device
            xyz = torch.empty((n, 3), dtype=torch.float64, device=dev)
            rgb = torch.empty((n, 3), dtype=torch.uint8, device=dev) if self.has_rgb else None
            sites = torch.empty((n, 2), dtype=torch.int32, device=dev)
            cam = N.ccam(self.camera)
            N.check(N.load().svb_cloud_export(N.ptr(self.records), 2, n, C.byref(cam), N.ptr(xyz), None, N.ptr(rgb),
                                              N.ptr(sites), None, None, N.stream_ptr()), "svb_cloud_export")
            N.count(1 if n else 0)
            self._host = (xyz.cpu().numpy(), rgb.cpu().numpy() if rgb is not None else None, sites.cpu().numpy())
        return self._host

    @property
    def xyz(self):
        return self._export()[0]

    @property
    def rgb(self):
        return self._export()[1]

    @property
    def sites(self):
        """int32 [n,2] pixel sites as (y, x)."""
        return self._export()[2]

    def _row(self, i):
        xyz, rgb, _ = self._export()
        if rgb is None:
            return [xyz[i, 0], xyz[i, 1], xyz[i, 2]]
        return [xyz[i, 0], xyz[i, 1], xyz[i, 2], rgb[i, 0], rgb[i, 1], rgb[i, 2]]

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self._row(k) for k in range(*i.indices(self.n))]
        if i < 0:
            i += self.n
        if not 0 <= i < self.n:
            raise IndexError("point index out of range")
        return self._row(i)

    def __iter__(self):
        for i in range(self.n):
            yield self._row(i)

    def __array__(self, dtype=None, copy=None):
        xyz, rgb, _ = self._export()
        a = xyz if rgb is None else np.concatenate([xyz, rgb.astype(np.float64)], axis=1)
        return a.astype(dtype) if dtype is not None else a


class DeviceColumn:
    """float64 (n,1) column living on the GPU (calculatePointErrors, functions.py:300-312)."""

    def __init__(self, t):
        self.t = t                        # float64 CUDA tensor [n]
        self._host = None

    def __len__(self):
        return int(self.t.numel())

    @property
    def shape(self):
        return (len(self), 1)

    def __array__(self, dtype=None, copy=None):
        if self._host is None:
            self._host = self.t.cpu().numpy().reshape(-1, 1)
        return self._host.astype(dtype) if dtype is not None else self._host

    def __getitem__(self, i):
        return np.asarray(self)[i]

    def __iter__(self):
        return iter(np.asarray(self))


class HueHistogram(dict):
    """dict key -> count exactly as calculateColourHistogram builds it (functions.py:215-226), plus the
    1024-bin device tensor the filter kernel reads.  key of bin b = str(round(np.float64(b/1000), 3))."""

    def __init__(self, bins_dev):
        bins = bins_dev.cpu().numpy().astype(np.int64)[:N.HUE_BINS]
        super().__init__({bin_key(b): int(c) for b, c in enumerate(bins) if c})
        self.bins_dev = bins_dev


def bin_key(b: int) -> str:
    return str(np.float64(b) / 1000.0)


def bins_from_dict(hist: dict, device):
    """A plain {key: count} dict (e.g. built by the reference) -> int32 [1024] CUDA tensor."""
    if isinstance(hist, HueHistogram):
        return hist.bins_dev
    bins = np.zeros(N.HIST_STRIDE, dtype=np.int32)
    for k, v in hist.items():
        b = int(round(float(k) * 1000.0))
        if not 0 <= b <= 1000:
            raise KeyError(k)
        bins[b] = int(v)
    return torch.from_numpy(bins).to(device)


class ImagePoints:
    """[x, y] float64 pairs of project3DPointsTo2DImagePoints (functions.py:201-209)."""

    def __init__(self, xy_dev):
        self.xy_dev = xy_dev              # float64 CUDA [n,2]
        self._host = None

    def __len__(self):
        return int(self.xy_dev.shape[0])

    def __array__(self, dtype=None, copy=None):
        if self._host is None:
            self._host = self.xy_dev.cpu().numpy()
        if dtype is None:
            return self._host
        return self._host.astype(dtype)       # float64 -> int32 truncates toward zero (stereovision.py:112)

    def __getitem__(self, i):
        r = np.asarray(self)[i]
        return [r[0], r[1]] if r.ndim == 1 else r

    def __iter__(self):
        for r in np.asarray(self):
            yield [r[0], r[1]]
