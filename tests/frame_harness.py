"""TEST HARNESS -- the reference's frame orchestrator (stereovision.py:26-247) restated on the drop-in helpers, without its
display layer.  The product story is INTEGRATION.md section 1: the UNMODIFIED orchestrator with `functions` routed to the
GPU helpers (stereo.vision_b200.dropin.route), which tests/test_gpu_frame.py::test_unmodified_orchestrator_on_gpu_helpers
executes.  This restatement exists so that every stage's product can be held to tests/golden/frames.npz (the recorded
products of the unmodified orchestrator): it returns a dict of the stage images instead of the tiled display image.
Exceptions are handled stage by stage exactly like the reference (a failing stage leaves a black image / no plane).
"""
import numpy as np

import stereo.vision_b200.functions as f

# stereovision.py:10-24
default_opts = {
    'crop_disparity': False, 'pause_playback': False, 'max_disparity': 128, 'ransac_trials': 600,
    'road_color_thresh': 10, 'loop': True, 'point_threshold': 0.05, 'image_tiles': True, 'img_size': (544, 1024),
    'threshold_option': 'previous', 'record_video': False, 'record_stats': False, 'video_filename': 'previous.avi'
}


def performStereoVision(imgL, imgR, prev_disp=None, opt=default_opts):
    """stereovision.py:26.  Returns (images, prev_disp, normal): `images` maps the reference's tile names
    ("Disparity", "Road Image", "Filtered Road Image", plus "Obstacles" / "Hull Mask") to arrays, and carries the
    frame's statistics (stereovision.py:32-37,101-125,201-202) under "stats"."""
    if 'frame' not in opt:                                                    # :27-30
        opt['frame'] = 1
    else:
        opt['frame'] += 1
    stats = {"Frame": opt['frame']}
    images = {}
    # 1. image processing (:44,46)
    imgL, imgR = f.preProcessImages(imgL, imgR)
    grayL, grayR = f.greyscale(imgL, imgR)
    # 2. disparity (:54-64)
    try:
        disparity = f.disparity(grayL, grayR, opt['max_disparity'], opt['crop_disparity'])
        disparity = f.disparityCleaning(disparity, opt['threshold_option'], prev_disp)
        prev_disp = disparity
    except Exception:
        print("Cannot compute the disparity.")
        disparity = f.getBlackImage()
    images["Disparity"] = disparity
    # 3. + 4. masks and clouds (:74-85)
    maskedDisparity = f.maskDisparity(disparity)
    cappedDisparity = f.capDisparity(disparity)
    points = f.projectDisparityTo3d(cappedDisparity, opt['max_disparity'], imgL)
    maskpoints = f.projectDisparityTo3d(maskedDisparity, opt['max_disparity'])
    # 5. plane (:92-125)
    planePoints = []
    normal = None
    try:
        normal, abc = f.RANSAC(maskpoints, opt['ransac_trials'])
        pointDifferences = f.calculatePointErrors(abc, points)
        points = f.computePlanarThreshold(points, pointDifferences, opt['point_threshold'])
        stats["Planar Points Before"] = len(points)
        histogram = f.calculateColourHistogram(points)
        points = f.filterPointsByHistogram(points, histogram, opt['road_color_thresh'])
        stats["Planar Points After"] = len(points)
        stats["Planar Pre-Filtering Accuracy"] = stats["Planar Points After"] / stats["Planar Points Before"]
        planePoints = f.project3DPointsTo2DImagePoints(points)
        planePoints = np.array(planePoints, np.int32).reshape((-1, 1, 2))
        stats["Computed Planar"] = 1
    except Exception as e:
        print("There was an error with generating a plane:", e)
        stats["Planar Points Before"] = stats["Planar Points After"] = stats["Planar Pre-Filtering Accuracy"] = "-"
        stats["Computed Planar"] = 0
    # 6. the overlay (:129-135) and the road image (:139-143)
    imageRoadMap = np.array(imgL, copy=True)
    for i in planePoints:
        imageRoadMap[i[0][1]][i[0][0]] = [0, 255, 0]
    images["Image Road Map"] = imageRoadMap
    try:
        roadImage = f.generatePointsAsImage(planePoints)
    except Exception as e:
        print("There was an error with generating a road image:", e)
        roadImage = f.getBlackImage()
    images["Road Image"] = roadImage
    # 7. cleaning (:155-161)
    try:
        cleanedRoadImage, resultingPoints = f.sanitiseRoadImage(roadImage, opt['img_size'])
    except Exception as e:
        print("There was an error with cleaning the road image:", e)
        cleanedRoadImage = f.getBlackImage()
        resultingPoints = planePoints
    images["Filtered Road Image"] = cleanedRoadImage
    # 8. obstacles (:170-189) and 9. hull centre (:196-206); the overlays onto the display image are left out
    try:
        roadHull, hullMask, obstacleImage = f.roadObstacles(cleanedRoadImage)
        images["Hull Mask"], images["Obstacles"], images["Road Hull"] = hullMask, obstacleImage, roadHull
        center = f.getCenterPoint(roadHull)
        stats["Center Point X"], stats["Center Point Y"] = center[0], center[1]
        images["Normal Line"] = f.getNormalVectorLine(center, normal, disparity)
    except Exception as e:
        print("There was an error in detecting obstacles / generating a hull:", e)
    images["stats"] = stats
    return images, prev_disp, normal
