"""Builds libloner_b200.so (hand-written sm_100a CUDA kernels + C-ABI) in-tree with nvcc.

    python -m loner_b200.build            # incremental (skips when the .so is newer than the sources)
    python -m loner_b200.build --force

The shared object lands in loner_b200/lib/ (git-ignored, but it travels to the GPU box with the
repo snapshot).  nvcc cross-compiles without a GPU.
"""
import glob
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libloner_b200.so")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
              "-I", INCLUDE]
OBJ_DIR = os.path.join(LIB_DIR, "obj")


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _deps():
    return sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(INCLUDE, "*.h"))


def up_to_date():
    if not os.path.exists(LIB_PATH):
        return False
    t = os.path.getmtime(LIB_PATH)
    return all(os.path.getmtime(s) <= t for s in _deps())


def build(force=False, verbose=False, test_switches=False):
    """test_switches: compile with -DLB_TEST_SWITCHES (fp32 CUDA-core MLP + LONER_B200_* environment switches for the
    cross-check tests and profiles/ experiments).  The default, shipped build has neither."""
    flag_file = os.path.join(LIB_DIR, "build_flags.txt")
    want = "test-switches" if test_switches else "shipped"
    have = open(flag_file).read().strip() if os.path.exists(flag_file) else None
    if have != want:
        force = True
    if not force and up_to_date():
        return LIB_PATH
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: cannot build libloner_b200.so")
    os.makedirs(OBJ_DIR, exist_ok=True)
    # one object per translation unit, compiled in parallel; a unit is rebuilt when it or any header is newer
    headers = glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(INCLUDE, "*.h"))
    t_hdr = max(os.path.getmtime(h) for h in headers)

    def compile_one(src):
        obj = os.path.join(OBJ_DIR, os.path.basename(src)[:-3] + ".o")
        if not force and os.path.exists(obj) and os.path.getmtime(obj) >= max(os.path.getmtime(src), t_hdr):
            return obj, ""
        cmd = [nvcc] + NVCC_FLAGS + (["-DLB_TEST_SWITCHES"] if test_switches else []) + \
            (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        return obj, r.stderr

    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        results = list(ex.map(compile_one, sources()))
    if verbose:
        print("".join(log for _, log in results))
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a"] + [o for o, _ in results] + ["-o", LIB_PATH, "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    with open(flag_file, "w") as f:
        f.write(want + "\n")
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, test_switches="--test-switches" in sys.argv))
