"""ORACLE / TEST INFRASTRUCTURE ONLY: stand-in for kornia.geometry (only calibration.undistort_points is restated)."""
from . import calibration  # noqa: F401
