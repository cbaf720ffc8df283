"""ORACLE -- TEST INFRASTRUCTURE ONLY.

CPU restatement (torch-CPU / numpy, fp32) of LONER's mapping hot path, used exclusively as the
checker for the CUDA path.  Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s
`cpu_baseline` / `--impl reference` legs may import anything from this package.  The product
package `loner_b200/` never does, and fails loudly when its CUDA library is missing.

Pinning status (see DESIGN.md "Oracle"):
  * sampling, compositing, target weights, loss, ray construction: PINNED against the reference's own
    Python (imported unmodified from /root/reference in the build container by
    oracle/reference_import.py; fixtures in tests/golden/ written by oracle/make_golden.py).
  * hash-grid encoding, fully-fused MLP, spherical harmonics (third-party tiny-cuda-nn 1.7, source
    absent from /root/reference, not installable here): **parity unpinned** -- restated from the
    published algorithm in oracle/tinycudann/__init__.py.
"""
