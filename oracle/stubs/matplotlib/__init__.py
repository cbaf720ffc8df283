"""ORACLE / TEST INFRASTRUCTURE ONLY: stand-in for matplotlib (reference optimizer.py:27-28 imports it for
visualize_loss, which is out of scope)."""
from _anything import Anything as _A
pyplot = _A()


def __getattr__(name):
    if name.startswith("__"):
        raise AttributeError(name)
    return _A()
