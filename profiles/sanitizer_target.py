"""Target of the compute-sanitizer runs (profiles/README.md, round 2): the smoke mapping step (all hot-path kernels:
ray build, sampler, fused sigma forward / backward, loss, Adam) and the fused-kernel unit tests with ragged tiles, in
ONE process.

    compute-sanitizer --tool memcheck  --log-file gpurun_out/r02_memcheck.txt  python profiles/sanitizer_target.py
    compute-sanitizer --tool racecheck --log-file gpurun_out/r02_racecheck.txt python profiles/sanitizer_target.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.chdir(ROOT)
import __graft_entry__ as G  # noqa: E402

G.smoke()
import pytest  # noqa: E402
rc = pytest.main(["-q", "-x", "-p", "no:cacheprovider", "tests/test_kernels_gpu.py", "-k",
                  "fused_sigma_forward_matches_two_kernel_path or test_sigma_fwd_bwd_golden or test_composite_loss_golden"])
sys.exit(int(rc))
