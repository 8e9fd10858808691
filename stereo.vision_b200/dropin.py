"""INTEGRATION.md section 1 as code: the reference's own `functions` module with the hot path routed to the GPU.

    import functions as f_ref
    import stereo.vision_b200.dropin as dropin
    f = dropin.route(f_ref)              # stereovision.py:3 then uses `f` unchanged

Helpers of the path (and of the SURVEY 8f rows built so far) resolve to stereo.vision_b200.functions; everything
else -- drawing, tiling, key handling, file I/O -- to the reference module.  Module constants (camera, masks) are
read from the reference module once, so both sides see the same values.
"""
from . import functions as f_gpu

HOT = ("disparityCleaning", "fillDisparity", "maskDisparity", "capDisparity", "projectDisparityTo3d",
       "RANSAC", "calculatePointErrors", "computePlanarThreshold", "calculateColourHistogram",
       "filterPointsByHistogram", "project3DPointsTo2DImagePoints", "generatePointsAsImage", "getBlackImage",
       "sanitiseRoadImage",                                  # stereovision.py:156, the first consumer of the road image
       "gammaChange", "preProcessImages", "greyscale",       # stereovision.py:44,46, the front end
       "fillAltDisparity", "getNormalVectorLine",            # the 'mean' option and the normal glyph
       "disparity",                                          # stereovision.py:56, the matcher stage
       "getCenterPoint")                                     # stereovision.py:200


class Routed:
    def __init__(self, f_ref, hot=HOT):
        self._ref, self._hot = f_ref, frozenset(hot)
        for name in ("carmask", "blackImg", "road_threshold_mask"):      # functions.py:29-35
            v = getattr(f_ref, name, None)
            if v is not None:
                setattr(f_gpu, name, v)
        for name in ("camera_focal_length_px", "stereo_camera_baseline_m", "image_centre_h", "image_centre_w"):
            if hasattr(f_ref, name):
                setattr(f_gpu, name, getattr(f_ref, name))

    def __getattr__(self, name):
        return getattr(f_gpu if name in self._hot else self._ref, name)


def route(f_ref, hot=HOT):
    return Routed(f_ref, hot)
