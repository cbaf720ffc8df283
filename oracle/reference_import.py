"""ORACLE / TEST INFRASTRUCTURE ONLY -- never imported by the product path (loner_b200/).

Imports the reference's own Python **unmodified** behind:
  * a `tinycudann` module: by default oracle/tinycudann (fp32 torch restatement of the tcnn modules, parity unpinned,
    see there) for CPU runs; tests/test_reference_dropin_gpu.py passes `loner_b200.tcnn` instead (the drop-in)
  * oracle/stubs/*  -- empty stand-ins for viz/IO imports + tiny pytorch3d/attrdict restatements

The reference tree is looked up at $LONER_REFERENCE_ROOT, then /root/reference (build container only), then
oracle/_ref (the verbatim copy oracle/make_ref_copy.py makes; git-ignored, travels to the GPU box).

Used by oracle/make_golden*.py (write tests/golden/*.pt), bench.py's CPU arm and the reference-facing tests.
"""
import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_PACKAGES = ("models", "mapping", "common", "src")


def reference_root():
    for p in (os.environ.get("LONER_REFERENCE_ROOT"), "/root/reference", os.path.join(_HERE, "_ref")):
        if p and os.path.isdir(os.path.join(p, "src", "models")):
            return p
    return None


REFERENCE_ROOT = reference_root() or "/root/reference"


def reference_available():
    return reference_root() is not None


def setup_paths():
    """Put shim + stubs + reference/src on sys.path (idempotent)."""
    root = reference_root()
    if root is None:
        raise RuntimeError("reference tree not found (neither /root/reference nor oracle/_ref: run "
                           "`python -m oracle.make_ref_copy` in the build container)")
    for p in (os.path.join(root, "src"), root, os.path.join(_HERE, "stubs"), _HERE):
        if p in sys.path:
            sys.path.remove(p)
        sys.path.insert(0, p)
    return root


def _purge():
    for name in list(sys.modules):
        if name.split(".")[0] in _REF_PACKAGES:
            m = sys.modules[name]
            f = getattr(m, "__file__", None) or ""
            if "/src/" in f or f.endswith("/src/__init__.py") or getattr(m, "__path__", None):
                del sys.modules[name]


def import_reference(tcnn=None):
    """Returns a namespace with the reference modules on the mapping hot path.  `tcnn` = the module to serve as
    `tinycudann` (default: the oracle's CPU shim).  Switching the binding re-imports the reference modules."""
    root = setup_paths()
    import types
    if tcnn is None:
        cur = sys.modules.get("tinycudann")
        if cur is not None and not os.path.abspath(getattr(cur, "__file__", None) or "").startswith(_HERE):
            del sys.modules["tinycudann"]            # another binding (e.g. loner_b200.tcnn) was installed earlier
        import tinycudann  # noqa: F401  (the oracle shim, found via oracle/ on sys.path)
        assert os.path.dirname(os.path.abspath(tinycudann.__file__)).startswith(_HERE), "wrong tinycudann picked up"
        tcnn = tinycudann
    bound = sys.modules.get("models.nerf_tcnn")
    if bound is not None and getattr(bound, "tcnn", None) is not tcnn:
        _purge()
    sys.modules["tinycudann"] = tcnn
    from models import model_tcnn, nerf_tcnn, rendering_tcnn, ray_sampling, losses
    from mapping import optimizer as ref_optimizer
    from common import pose_utils, ray_utils, settings, pose, sensors, frame
    from mapping import keyframe
    assert nerf_tcnn.tcnn is tcnn
    assert os.path.abspath(nerf_tcnn.__file__).startswith(os.path.abspath(root)), nerf_tcnn.__file__
    ns = types.SimpleNamespace(
        model_tcnn=model_tcnn, nerf_tcnn=nerf_tcnn, rendering_tcnn=rendering_tcnn,
        ray_sampling=ray_sampling, losses=losses, optimizer=ref_optimizer,
        pose_utils=pose_utils, ray_utils=ray_utils, settings=settings, pose=pose,
        sensors=sensors, frame=frame, keyframe=keyframe, tinycudann=tcnn, root=root)
    return ns
