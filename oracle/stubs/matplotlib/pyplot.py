"""ORACLE / TEST INFRASTRUCTURE ONLY."""
from _anything import Anything as _A


def __getattr__(name):
    if name.startswith("__"):
        raise AttributeError(name)
    return _A()
