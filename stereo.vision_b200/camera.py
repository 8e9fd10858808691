"""Camera constants of the path.

The reference keeps these as module globals (functions.py:15-24).  They are grouped in a small
immutable record here so that the stress configuration (2048x1088) can carry
a scaled camera next to the default one; the drop-in module re-exports the default values under the
reference's own names.
"""
from dataclasses import dataclass


@dataclass(frozen=True)
class Camera:
    focal_px: float = 399.9745178222656      # functions.py:15
    baseline_m: float = 0.2090607502         # functions.py:19
    centre_h: float = 262.0                  # functions.py:21
    centre_w: float = 474.5                  # functions.py:22

    def scaled(self, s: float) -> "Camera":
        """Camera of the same rig imaged at s times the resolution."""
        return Camera(self.focal_px * s, self.baseline_m, self.centre_h * s, self.centre_w * s)


DEFAULT_CAMERA = Camera()
MAX_DISPARITY = 128                          # functions.py:24 (passed around, never used on the path)
IMAGE_SHAPE = (544, 1024)                    # stereovision.py:19
