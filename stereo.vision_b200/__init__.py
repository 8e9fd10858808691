"""stereo.vision_b200 -- B200-native hot path of thien/stereo.vision."""
from .camera import Camera, DEFAULT_CAMERA, MAX_DISPARITY, IMAGE_SHAPE  # noqa: F401
