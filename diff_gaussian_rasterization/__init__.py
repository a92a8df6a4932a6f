"""Drop-in shim: `from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer`
(python_ordinary/scripts/splatam.py:39, python_ordinary/utils/recon_helpers.py:2) resolves to the
B200-native implementation in isogs-slam-cpp_b200/."""
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
import isogs_b200  # noqa: E402

GaussianRasterizationSettings = isogs_b200.pkg.GaussianRasterizationSettings
GaussianRasterizer = isogs_b200.pkg.GaussianRasterizer
__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer"]
