import functools, inspect


class _Cfg(dict):
    __getattr__ = dict.get


class ConfigMixin:
    config_name = "config.json"

    @property
    def config(self):
        return getattr(self, "_vc_config", _Cfg())

    @classmethod
    def from_config(cls, config, **kw):
        sig = inspect.signature(cls.__init__).parameters
        return cls(**{k: v for k, v in dict(config, **kw).items() if k in sig})


def register_to_config(init):
    @functools.wraps(init)
    def wrapper(self, *a, **kw):
        sig = inspect.signature(init)
        bound = sig.bind(self, *a, **kw)
        bound.apply_defaults()
        cfg = _Cfg({k: v for k, v in bound.arguments.items() if k != "self"})
        init(self, *a, **kw)
        prev = getattr(self, "_vc_config", None)
        if prev is not None:
            merged = _Cfg(prev); merged.update(cfg); cfg = merged
        object.__setattr__(self, "_vc_config", cfg)
    return wrapper
