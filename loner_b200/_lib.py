"""ctypes binding of libloner_b200.so (C-ABI declared in include/loner_b200.h).

There is no CPU fallback and no alternative backend: if the shared library is missing or a call
fails, this module raises.  `load()` is lazy so that host-only code (config, synthetic scenes,
ray construction) can be imported on a machine without the built library.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libloner_b200.so")

_lib = None


class GridConfig(C.Structure):
    _fields_ = [("n_levels", C.c_int32), ("n_features_per_level", C.c_int32),
                ("log2_hashmap_size", C.c_int32), ("base_resolution", C.c_int32),
                ("per_level_scale", C.c_float), ("mlp_width", C.c_int32), ("mlp_hidden_layers", C.c_int32)]


class LossConfig(C.Structure):
    _fields_ = [("kind", C.c_int32), ("scale_factor", C.c_float), ("min_js_score", C.c_float),
                ("max_js_score", C.c_float), ("js_alpha", C.c_float), ("min_depth_eps", C.c_float),
                ("depth_eps", C.c_float), ("los_lambda", C.c_float), ("depthloss_lambda", C.c_float),
                ("n_opaque", C.c_float), ("n_total_samples", C.c_float)]


LOSS_KINDS = {"L1_JS": 0, "L2_JS": 1, "L1_LOS": 2, "L2_LOS": 3}

_P = C.c_void_p
_I = C.c_int
_F = C.c_float
_L = C.c_int64

# name -> (restype, argtypes); must list every symbol include/loner_b200.h declares
SIGNATURES = {
    "lb_version": (_I, []),
    "lb_last_error": (C.c_char_p, []),
    "lb_plan_create": (_I, [C.POINTER(GridConfig), C.POINTER(_P)]),
    "lb_plan_destroy": (None, [_P]),
    "lb_plan_set_l2_bytes": (_I, [_P, _L]),
    "lb_build_flags": (_I, []),
    "lb_plan_param_counts": (_I, [_P, C.POINTER(_L), C.POINTER(_L)]),
    "lb_plan_level": (_I, [_P, _I, C.POINTER(_F), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32),
                           C.POINTER(C.c_uint32), C.POINTER(C.c_int32)]),
    "lb_pose_to_rot_origin": (_I, [_P, _P, _F, _I, _P, _P, _P]),
    "lb_pose_to_rot_origin_bwd": (_I, [_P, _F, _I, _P, _P, _P, _P]),
    "lb_build_lidar_rays": (_I, [_P, _P, _P, _I, _I, _P, _P, _F, _F, _F, _F, _F, _P, _P, _P]),
    "lb_voxel_splat_max": (_I, [_P, _P, _P, _P, _I, _I, _P, _I, _P, _I, _P, _I, _P, _P]),
    "lb_build_camera_rays": (_I, [_P, _P, _P, _P, _I, _P, _I, _I, _P, _P, _F, _P, _P, _P]),
    "lb_build_lidar_rays_bwd": (_I, [_P, _P, _I, _I, _P, _P, _F, _P, _P, _P, _P]),
    "lb_sample_occ": (_I, [_P, _P, _I, _P, _P, _P, _I, _I, _F, _P, _P, _P, _P]),
    "lb_sample_uniform": (_I, [_P, _P, _P, _I, _I, _F, _P, _P]),
    "lb_encode_fwd": (_I, [_P, _P, _P, _L, _P, _P]),
    "lb_encode_bwd": (_I, [_P, _P, _P, _P, _L, _P, _P, _P]),
    "lb_sigma_fwd_enc_optional": (_I, [_P]),
    "lb_sigma_fwd": (_I, [_P, _P, _P, _P, _P, _I, _I, _P, _P, _P, _P]),
    "lb_sigma_fwd_noalloc": (_I, [_P, _P, _P, _P, _P, _I, _I, _P, _P]),
    "lb_sigma_fwd_jvp_supported": (_I, [_P]),
    "lb_sigma_fwd_jvp": (_I, [_P, _P, _P, _P, _P, _I, _I, _P, _P, _P, _P, _P]),
    "lb_rays_grad_from_jvp": (_I, [_P, _P, _P, _I, _I, _P, _P]),
    "lb_sigma_bwd": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _P, _P, _P, _P]),
    "lb_composite_fwd": (_I, [_P, _P, _P, _P, _I, _I, _I, _P, _P, _P, _P, _P]),
    "lb_composite_bwd": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _P, _P, _P]),
    "lb_composite_loss": (_I, [_P, _P, _P, _P, _P, C.POINTER(LossConfig), _P, _I, _I,
                               _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "lb_count_opaque": (_I, [_P, _P, _I, _P, _P]),
    "lb_adam_step": (_I, [_P, _P, _P, _P, _L, C.c_double, C.c_double, C.c_double, C.c_double, _I, _F, _I, _P]),
    "lb_occ_grad": (_I, [_P, _P, _P, _F, _I, _I, _I, _P, _P]),
    "lb_sgd_step": (_I, [_P, _P, _L, _F, _I, _P]),
    "lb_sh4_fwd": (_I, [_P, _L, _P, _P]),
    "lb_sh4_bwd": (_I, [_P, _P, _L, _P, _P]),
    "lb_net_fwd": (_I, [_P, _L, _I, _I, _I, _I, _P, _P, _P, _P]),
    "lb_net_bwd": (_I, [_P, _L, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P]),
    "lb_composite_rgb_fwd": (_I, [_P, _P, _I, _I, _I, _I, _P, _P]),
    "lb_composite_rgb_bwd": (_I, [_P, _P, _P, _I, _I, _I, _I, _P, _P, _P]),
}


class LonerB200Error(RuntimeError):
    pass


def load():
    """Loads the shared library (once). Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise LonerB200Error(
            f"{LIB_PATH} is missing: build it with `python -m loner_b200.build` (needs nvcc). "
            "loner_b200 has no CPU or PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, what=""):
    if rc != 0:
        msg = load().lb_last_error().decode("utf-8", "replace")
        raise LonerB200Error(f"{what} failed with status {rc}: {msg}")


# Launch accounting: every compute entry point launches exactly one kernel of ours.  `LAUNCHES` counts
# them; with `TIMING` on, each call is bracketed by CUDA events on the launching stream (bench.py reads
# the per-kernel device time for the roofline figures).
LAUNCHES = {}
TIMING = {"on": False, "events": []}
_HOST_ONLY = {"lb_plan_create", "lb_plan_param_counts", "lb_plan_level", "lb_plan_set_l2_bytes"}


def test_switches():
    """True for a library built with -DLB_TEST_SWITCHES (fp32 CUDA-core MLP cross-check available)."""
    return bool(load().lb_build_flags() & 1)


def call(name, *args):
    """Calls an int-returning entry point and raises on a non-zero status."""
    fn = getattr(load(), name)
    if name in _HOST_ONLY:
        check(fn(*args), name)
        return
    LAUNCHES[name] = LAUNCHES.get(name, 0) + 1
    if TIMING["on"]:
        import torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        check(fn(*args), name)
        e1.record()
        TIMING["events"].append((name, e0, e1))
    else:
        check(fn(*args), name)


def reset_counters():
    LAUNCHES.clear()
    TIMING["events"].clear()


def kernel_times():
    """{name: (launches, total_ms)} for calls made while TIMING was on (synchronises)."""
    import torch
    torch.cuda.synchronize()
    out = {}
    for name, e0, e1 in TIMING["events"]:
        n, t = out.get(name, (0, 0.0))
        out[name] = (n + 1, t + e0.elapsed_time(e1))
    return out
