import logging as _pylog


def is_torch_version(op, ver):
    return True


class logging:  # noqa: N801 - mirrors diffusers.utils.logging
    @staticmethod
    def get_logger(name):
        return _pylog.getLogger(name)
