"""isogs-slam-cpp_b200: B200-native (sm_100a) differentiable Gaussian rasterizer and SLAM step
kernels behind a C ABI (include/isogs_rast.h).  The directory name is not a Python identifier;
import it through the repo-root helper `isogs_b200` (module name `isogs_slam_cpp_b200`) or through
the drop-in shim `diff_gaussian_rasterization`."""
from . import _lib, raster  # noqa: F401
from ._lib import build, lib, make_camera, make_pose  # noqa: F401
from .rasterizer import GaussianRasterizationSettings, GaussianRasterizer, camera_from_settings  # noqa: F401
