"""Stand-in for VideoX-Fun's TeaCache (un-vendored; restated from its published behaviour, parity unpinned for the class
itself).  It carries exactly the attributes the reference's code touches — wan_transformer3d.py:205-245 (skip decision),
wan_transformer3d_versecrafter.py:384-417, :438-441 (residual reuse, cnt / reset), wan_transformer3d.py:828-847
(enable / share / disable) — so that the reference's own TeaCache BRANCHES, which are in the tree, can run in the parity
tests (tests/test_pipeline_boundary.py).  TEST INFRASTRUCTURE: not imported by the product path, which has its own class."""
import numpy as np
import torch


class TeaCache:
    def __init__(self, coefficients, num_steps, rel_l1_thresh=0.0, num_skip_start_steps=0, offload=True):
        self.coefficients = coefficients
        self.num_steps = num_steps
        self.rel_l1_thresh = rel_l1_thresh
        self.num_skip_start_steps = num_skip_start_steps
        self.offload = offload
        self.rescale_func = np.poly1d(coefficients)
        self.cnt = 0
        self.should_calc = True
        self.accumulated_rel_l1_distance = 0
        self.previous_modulated_input = None
        self.previous_residual = None
        self.previous_residual_cond = None
        self.previous_residual_uncond = None

    @staticmethod
    def compute_rel_l1_distance(prev, cur):
        rel = torch.abs(prev - cur).mean() / torch.abs(prev).mean()
        return rel.cpu().item()

    def reset(self):
        self.cnt = 0
        self.should_calc = True
        self.accumulated_rel_l1_distance = 0
        self.previous_modulated_input = None
        self.previous_residual = None
        self.previous_residual_cond = None
        self.previous_residual_uncond = None
