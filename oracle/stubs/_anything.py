"""ORACLE / TEST INFRASTRUCTURE ONLY. Placeholder object for viz/IO libraries the reference imports
at module scope (torchviz, matplotlib, open3d, kornia) but that the mapping hot path never calls."""


class Anything:
    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        return Anything()

    def __call__(self, *a, **k):
        raise RuntimeError("oracle stub: this third-party call is outside the mapping hot path")
