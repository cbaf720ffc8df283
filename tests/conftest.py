import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "reference: needs /root/reference (build container only)")


def pytest_collection_modifyitems(config, items):
    """`gpu` items are skipped (not failed) on a box without a CUDA device, `reference` items where /root/reference
    is absent -- so a plain `pytest tests/` is green in the build container and on the GPU box alike."""
    import torch
    has_gpu = torch.cuda.is_available()
    has_ref = os.path.isdir("/root/reference/src")
    for item in items:
        if "gpu" in item.keywords and not has_gpu:
            item.add_marker(pytest.mark.skip(reason="needs a CUDA device"))
        if "reference" in item.keywords and not has_ref:
            item.add_marker(pytest.mark.skip(reason="needs /root/reference (build container only)"))


def load_golden(name):
    import torch
    return torch.load(os.path.join(GOLDEN_DIR, f"{name}.pt"), weights_only=False)


@pytest.fixture(scope="session")
def golden():
    return load_golden


def golden_params(g):
    """Re-create the flat sigma parameter vector a golden fixture was generated with."""
    from oracle import loner_oracle as O
    p = O.init_sigma_params(g["log2_T"], g["param_seed"])
    p[O.MLP_PARAMS:] *= g["table_scale"]
    assert abs(p.double().sum().item() - g["params_checksum"]) < 1e-6 * max(1.0, abs(g["params_checksum"]))
    return p


def golden_rgb_params(g):
    """(sigma params, intensity table, intensity MLP) an rgb golden fixture was generated with."""
    from oracle import loner_oracle as O
    p = golden_params(g)
    table, mlp = O.init_intensity_params(g["log2_T_int"], 3, g["param_seed"])
    table = table * g["table_scale"]
    assert abs(table.double().sum().item() - g["table_int_checksum"]) < 1e-6 * max(1.0, abs(g["table_int_checksum"]))
    assert abs(mlp.double().sum().item() - g["mlp_int_checksum"]) < 1e-6 * max(1.0, abs(g["mlp_int_checksum"]))
    return p, table, mlp


def c2_shape_random(g):
    """Per-chunk (u1, u2, noise) the reference consumed while tests/golden/c2_shape_ogm_l1js.pt was recorded:
    oracle/portable_rng.py streams (torch.rand call k -> stream 2k, torch.randn call k -> stream 2k + 1)."""
    from oracle import portable_rng as R
    n, S, chunk, base = g["rays"].shape[0], g["n_samples"], g["chunk"], g["rng_base_stream"]
    out = []
    for c, i in enumerate(range(0, n, chunk)):
        m = min(chunk, n - i)
        out.append((R.uniform((m, S // 2), base + 2 * (2 * c)), R.uniform((m, S // 2), base + 2 * (2 * c + 1)),
                    R.normal((m, S), base + 2 * c + 1)))
    return out


def c2_shape_params(g, state):
    """Flat sigma parameters of a state of the benchmark-shaped fixture (keys 2000, 1, "l2js_2000")."""
    from oracle import loner_oracle as O
    st = g["states"][state]
    p = O.init_sigma_params(g["log2_T"], g["param_seed"])
    p[O.MLP_PARAMS:] *= float(st["table_scale"])
    assert abs(p.double().sum().item() - st["params_checksum"]) < 1e-6 * max(1.0, abs(st["params_checksum"]))
    return p
