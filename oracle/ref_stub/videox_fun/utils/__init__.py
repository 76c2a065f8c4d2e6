"""Stand-in for videox_fun.utils (un-vendored).  `cfg_skip` is restated from VideoX-Fun's published decorator (SURVEY.md
App. A.11; parity unpinned for the decorator itself): once `current_steps >= num_inference_steps * (1 - cfg_skip_ratio)` a
batch of >= 2 is cut to its second half (the text-conditioned samples), every tensor / list / tuple argument with it, and the
result is returned twice so that the caller's `.chunk(2)` still works.  Inactive while `cfg_skip_ratio` is None (the default),
which is every parity run except the cfg-skip case of tests/test_pipeline_boundary.py.  TEST INFRASTRUCTURE."""
import functools

import numpy as np
import torch


def cfg_skip():
    def deco(fn):
        @functools.wraps(fn)
        def wrapper(self, x, *args, **kwargs):
            bs = len(x)
            ratio = getattr(self, "cfg_skip_ratio", None)
            if bs >= 2 and ratio is not None and self.current_steps >= self.num_inference_steps * (1 - ratio):
                half = bs // 2
                cut = lambda a: a[half:] if isinstance(a, (torch.Tensor, list, tuple, np.ndarray)) else a
                out = fn(self, x[half:], *[cut(a) for a in args], **{k: cut(v) for k, v in kwargs.items()})
                return torch.cat([out, out], dim=0)
            return fn(self, x, *args, **kwargs)
        return wrapper
    return deco
