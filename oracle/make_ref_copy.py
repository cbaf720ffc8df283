"""ORACLE / TEST INFRASTRUCTURE ONLY.  Container-only recipe (needs /root/reference).

Copies the reference's own Python for the mapping hot path VERBATIM into oracle/_ref/ (git-ignored: the sources never
enter this repository's history; NOT gpurun-ignored, so the copy travels to the GPU box like a built .so), so that

  * `bench.py --impl reference` and the `cpu_baseline` leg time the reference's own code on the GPU box's host cores
    (behind oracle/tinycudann, the fp32 CPU restatement of its tiny-cuda-nn dependency), and
  * tests/test_reference_dropin_gpu.py runs the UNMODIFIED reference (models/model_tcnn.Model, Optimizer.compute_loss)
    on CUDA on top of `loner_b200.tcnn` -- the drop-in INTEGRATION.md section 1 describes.

    python -m oracle.make_ref_copy          # also run by __graft_entry__.build() when /root/reference is present

Nothing under loner_b200/ may import oracle/_ref (tests/test_abi.py checks the product for `oracle` imports).
"""
import filecmp
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC_ROOT = os.environ.get("LONER_REFERENCE_SRC", "/root/reference")
DST_ROOT = os.path.join(HERE, "_ref")
# the packages on the path and what they import from each other
FILES = ["src/__init__.py"] + \
    [f"src/models/{f}" for f in ("__init__.py", "losses.py", "model_tcnn.py", "nerf_tcnn.py", "ray_sampling.py", "rendering_tcnn.py")] + \
    [f"src/mapping/{f}" for f in ("__init__.py", "keyframe.py", "keyframe_manager.py", "optimizer.py")] + \
    [f"src/common/{f}" for f in ("__init__.py", "frame.py", "pose.py", "pose_utils.py", "ray_utils.py", "sensors.py", "settings.py")]


def available():
    return os.path.isdir(os.path.join(SRC_ROOT, "src", "models"))


def copied():
    return all(os.path.exists(os.path.join(DST_ROOT, f)) for f in FILES)


def main():
    if not available():
        print(f"make_ref_copy: {SRC_ROOT} not found (GPU box?) -- keeping the existing copy: {copied()}")
        return copied()
    n = 0
    for f in FILES:
        src, dst = os.path.join(SRC_ROOT, f), os.path.join(DST_ROOT, f)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if not os.path.exists(dst) or not filecmp.cmp(src, dst, shallow=False):
            shutil.copyfile(src, dst)
            n += 1
    print(f"make_ref_copy: {len(FILES)} files under {DST_ROOT} ({n} refreshed)")
    return True


if __name__ == "__main__":
    sys.exit(0 if main() else 1)
