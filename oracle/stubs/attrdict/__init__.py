"""ORACLE / TEST INFRASTRUCTURE ONLY: minimal `attrdict.AttrDict` (reference src/common/settings.py:19,51).
Attribute access on a dict; nested dicts are wrapped on read, like the original package."""


class AttrDict(dict):
    def __getattr__(self, name):
        try:
            v = self[name]
        except KeyError:
            raise AttributeError(name)
        if isinstance(v, dict) and not isinstance(v, AttrDict):
            v = type(self)(v) if type(self) is AttrDict else AttrDict(v)
        return v

    def __setattr__(self, name, value):
        self[name] = value
