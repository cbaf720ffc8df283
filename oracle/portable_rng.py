"""ORACLE / TEST INFRASTRUCTURE ONLY.

Portable stand-ins for torch.rand / torch.randn: integer hashing in int64 tensor arithmetic followed by ONE exact
conversion to fp32, so that every machine and torch build reproduces the same bits from (stream, shape) alone and
large fixtures need not store the random tensors the reference consumed.  oracle/make_golden_c2.py patches
torch.rand / torch.randn with these for the duration of the reference's compute_loss; the tests replay them.

    uniform(shape, stream)  k / 2^24, k in [0, 2^24): exactly representable, in [0, 1)
    normal(shape, stream)   Irwin-Hall: (sum of twelve 16-bit hashes + 6) / 2^16 - 6: mean 0, variance 1, exact in fp32
"""
import math

import torch

_M = 0xFFFFFFFF


def _mix(x):
    x = x ^ (x >> 16)
    x = (x * 0x45D9F3B) & _M
    x = x ^ (x >> 16)
    x = (x * 0x45D9F3B) & _M
    return x ^ (x >> 16)


def _hash(n, stream, salt):
    i = torch.arange(n, dtype=torch.int64)
    return _mix((i * 2654435761 + (stream * 40503 + salt * 9176 + 12345)) & _M)


def uniform(shape, stream):
    n = int(math.prod(shape))
    return ((_hash(n, stream, 0) >> 8).to(torch.float32) / 16777216.0).reshape(tuple(shape))


def normal(shape, stream):
    n = int(math.prod(shape))
    acc = torch.zeros(n, dtype=torch.int64)
    for k in range(12):
        acc += _hash(n, stream, 101 + k) & 0xFFFF
    return ((acc + 6 - 6 * 65536).to(torch.float32) / 65536.0).reshape(tuple(shape))


class Patch:
    """Context manager: torch.rand -> uniform(shape, stream = 2 * call), torch.randn -> normal(shape, 2 * call + 1);
    results are moved to the `device=` the caller asked for and logged (`.rand`, `.randn`)."""

    def __init__(self, base_stream=0):
        self.base, self.rand, self.randn = base_stream, [], []

    @staticmethod
    def _shape(args):
        return tuple(args[0]) if len(args) == 1 and not isinstance(args[0], int) else tuple(args)

    def __enter__(self):
        self._orig = (torch.rand, torch.randn)

        def fake_rand(*a, **k):
            t = uniform(self._shape(a), self.base + 2 * len(self.rand))
            self.rand.append(t)
            return t.to(k.get("device", "cpu"))

        def fake_randn(*a, **k):
            t = normal(self._shape(a), self.base + 2 * len(self.randn) + 1)
            self.randn.append(t)
            return t.to(k.get("device", "cpu"))
        torch.rand, torch.randn = fake_rand, fake_randn
        return self

    def __exit__(self, *exc):
        torch.rand, torch.randn = self._orig
        return False
