"""ORACLE / TEST INFRASTRUCTURE ONLY: stand-in for kornia.geometry (camera undistortion; not on the LiDAR path)."""
from _anything import Anything as _A
calibration = _A()
