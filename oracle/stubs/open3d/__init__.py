"""ORACLE / TEST INFRASTRUCTURE ONLY: empty stand-in so the reference's module-scope imports succeed."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(__file__)))
from _anything import Anything as _A


def __getattr__(name):
    if name.startswith("__"):
        raise AttributeError(name)
    return _A()
