"""ORACLE / TEST INFRASTRUCTURE ONLY: torch restatement of the three pytorch3d.transforms functions the
reference's pose code uses (src/common/pose_utils.py:340,360,379; pose.py:106,174)."""
from . import transforms  # noqa: F401
