"""CPU tests of the drop-in boundary: the C-ABI library builds, loads, exports every symbol that
include/loner_b200.h declares, and its host-only entry points (plan / level table / error path) agree
with the oracle's restatement of the tcnn level layout.  No compute calls (no GPU here)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from loner_b200 import build, _lib
    build.build()
    return _lib.load()


def test_header_symbols_all_exported(lib):
    from loner_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "loner_b200.h")).read()
    declared = set(re.findall(r"LB_API[^;(]*?\b(lb_\w+)\s*\(", hdr))
    assert len(declared) >= 19
    assert declared == set(_lib.SIGNATURES), "ctypes binding and header disagree"
    for name in declared:
        assert hasattr(lib, name), f"{name} not exported"
    assert lib.lb_version() == 100


@pytest.mark.parametrize("log2T", [14, 18, 19, 22])
def test_plan_level_table_matches_oracle(lib, log2T):
    from loner_b200 import ops
    from oracle import tinycudann as T
    plan = ops.Plan(log2_hashmap_size=log2T)
    lv, entries = T.level_table(16, 16, log2T)
    assert plan.n_table_params == entries * 2
    assert plan.n_mlp_params == 3072
    for l in range(16):
        a, b = plan.level(l), lv[l]
        assert (a["res"], a["offset"], a["size"], a["dense"]) == (b["res"], b["offset"], b["size"], b["dense"])
        assert a["scale"] == b["scale"]


def test_errors_are_loud(lib):
    from loner_b200 import ops, _lib
    with pytest.raises(_lib.LonerB200Error, match="unsupported"):
        ops.Plan(n_features_per_level=4)
    with pytest.raises(_lib.LonerB200Error):
        ops.Plan(log2_hashmap_size=40)
    import torch
    with pytest.raises(_lib.LonerB200Error, match="CUDA tensor"):
        ops.sample_uniform(torch.zeros(4, 13), torch.zeros(4, 8), 8, 1.0)


def test_no_oracle_import_in_product():
    """The product package must never reach into oracle/ (or any CPU fallback)."""
    pkg = os.path.join(ROOT, "loner_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith(".py"):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), f"{f} imports oracle"
