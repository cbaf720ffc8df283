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
