"""``src/lib/data_generator/util.py`` of the reference, same function."""
import numpy as np


def camera_parameters(camera_config):
    """pose (camera-to-world 4x4) and 3x3 intrinsics of a camera config ->
    (E^-1 4x4 world-to-camera, K padded to 4x4)   (reference util.py:4-10)."""
    pose = camera_config['pose']
    k = np.reshape(camera_config['intrinsics'], (3, 3))
    k4 = np.eye(4, dtype=np.float64)
    k4[:3, :3] = k
    return np.linalg.inv(pose), k4
