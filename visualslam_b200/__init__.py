"""visualslam_b200 -- B200-native (sm_100a) KeyPointDetection hot path of JacobYoung115/VisualSLAM.

Product code: CUDA kernels + C ABI in csrc/ (libvslam_b200.so), C++17 host mirror of the reference surface in
host/, and this ctypes binding.  Nothing here imports oracle/.
"""
from .api import (COMPACT, CORNER8_DTYPE, CORNER_DTYPE, DEVICE, HOST, KEYPOINT_DTYPE, MAP_BASE, MAP_DOG, MAP_GAUSS, MAP_GRADX, MAP_GRADY,
                  MAP_MAG, MAP_ORIENT, Context, Pyramid, VslamError, calculate_sigma, expand_corners, gauss_taps, load)

__all__ = ["Context", "Pyramid", "VslamError", "load", "calculate_sigma", "gauss_taps", "KEYPOINT_DTYPE",
           "CORNER_DTYPE", "CORNER8_DTYPE", "COMPACT", "expand_corners", "HOST", "DEVICE", "MAP_BASE", "MAP_GAUSS", "MAP_DOG", "MAP_GRADX", "MAP_GRADY", "MAP_MAG",
           "MAP_ORIENT"]
