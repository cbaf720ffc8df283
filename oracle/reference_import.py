"""ORACLE / TEST INFRASTRUCTURE ONLY -- never imported by the product path (loner_b200/).

Imports the reference's own Python **unmodified** from /root/reference (read-only; exists only in
the build container, NOT on the GPU box) behind:
  * oracle/tinycudann  -- fp32 torch restatement of the tcnn modules (parity unpinned, see there)
  * oracle/stubs/*     -- empty stand-ins for viz/IO imports + tiny pytorch3d/attrdict restatements

Used by oracle/make_golden.py (writes tests/golden/*.pt) and by the container-only tests that pin
oracle/loner_oracle.py against the reference code itself.
"""
import os
import sys

REFERENCE_ROOT = os.environ.get("LONER_REFERENCE_ROOT", "/root/reference")
_HERE = os.path.dirname(os.path.abspath(__file__))


def reference_available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "src", "models"))


def setup_paths():
    """Put shim + stubs + reference/src on sys.path (idempotent)."""
    if not reference_available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT} (it does not travel to the GPU box)")
    for p in (os.path.join(REFERENCE_ROOT, "src"), REFERENCE_ROOT, os.path.join(_HERE, "stubs"), _HERE):
        if p not in sys.path:
            sys.path.insert(0, p)


def import_reference():
    """Returns a namespace with the reference modules on the mapping hot path."""
    setup_paths()
    import types
    import tinycudann  # noqa: F401  (the oracle shim, found via oracle/ on sys.path)
    assert os.path.dirname(os.path.abspath(tinycudann.__file__)).startswith(_HERE), "wrong tinycudann picked up"
    from models import model_tcnn, nerf_tcnn, rendering_tcnn, ray_sampling, losses
    from mapping import optimizer as ref_optimizer
    from common import pose_utils, ray_utils, settings, pose, sensors, frame
    from mapping import keyframe
    ns = types.SimpleNamespace(
        model_tcnn=model_tcnn, nerf_tcnn=nerf_tcnn, rendering_tcnn=rendering_tcnn,
        ray_sampling=ray_sampling, losses=losses, optimizer=ref_optimizer,
        pose_utils=pose_utils, ray_utils=ray_utils, settings=settings, pose=pose,
        sensors=sensors, frame=frame, keyframe=keyframe, tinycudann=tinycudann)
    return ns
