import logging as _pylog


def is_torch_version(op, ver):
    return True


class logging:  # noqa: N801 - mirrors diffusers.utils.logging
    @staticmethod
    def get_logger(name):
        return _pylog.getLogger(name)


class BaseOutput:
    """dataclass base of diffusers' pipeline outputs (only attribute access is used)."""


def replace_example_docstring(doc):
    def deco(fn):
        return fn
    return deco
