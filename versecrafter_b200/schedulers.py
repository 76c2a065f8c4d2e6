"""Samplers of the denoising loop as kernels of libvc_b200.so.

`FlowUniPCMultistepScheduler` keeps the surface the reference pipeline touches on its default sampler "Flow_Unipc"
(/root/reference/inference/versecrafter_inference.py:135,:254-262; pipeline_wan_versecrafter.py:750-752
`set_timesteps(n, device=, shift=)`, `.timesteps`, :868 `.order`, :909 `.step(noise_pred, t, latents, return_dict=False)`):
the upstream class lives in VideoX-Fun (`videox_fun.utils.fm_solvers_unipc`, un-vendored).  Per step the host computes
thirteen scalars in float64 and ONE kernel (`vc_cfg_unipc_step`) does the CFG combine, the x0 conversion, the UniPC
corrector and the predictor over the latents; the x0 history and the last corrected sample are fp32 device buffers.
`step_cfg()` is the fused entry the in-repo loop uses (CFG inside the kernel); `step()` is the pipeline-compatible one.
There is no CPU path: tensors must be CUDA bf16."""
import ctypes
import math

import numpy as np
import torch

from . import _lib, ops


def _lam(sigma):
    """log(alpha) - log(sigma) of the flow-matching path alpha = 1 - sigma; +inf at sigma = 0, -inf at sigma = 1."""
    if sigma <= 0.0:
        return math.inf
    if sigma >= 1.0:
        return -math.inf
    return math.log1p(-sigma) - math.log(sigma)


def _expm1(v):
    return -1.0 if v == -math.inf else math.expm1(v)


def _bh(h, order, rk):
    """(h_phi_1, B_h, rhos) of the predict-x0 bh2 update with step h in lambda and (for order 2) the ratio rk of the
    previous step; rhos solves R rho = b with R = [[1, 1], [rk, 1]] (corrector; the predictor's order-2 rho is 1/2)."""
    hh = -h
    h_phi_1 = _expm1(hh)
    B_h = h_phi_1
    if hh == -math.inf:
        return h_phi_1, B_h, None
    h_phi_k = h_phi_1 / hh - 1.0
    b1 = h_phi_k / B_h
    if order == 1:
        return h_phi_1, B_h, (0.5,)
    h_phi_k2 = h_phi_k / hh - 0.5
    b2 = h_phi_k2 * 2.0 / B_h
    # [1 1; rk 1] [r0; r1] = [b1; b2]
    r0 = (b1 - b2) / (1.0 - rk)
    return h_phi_1, B_h, (r0, b1 - r0)


class FlowUniPCMultistepScheduler:
    order = 1
    init_noise_sigma = 1.0

    def __init__(self, num_train_timesteps=1000, solver_order=2, prediction_type="flow_prediction", shift=1.0,
                 use_dynamic_shifting=False, predict_x0=True, solver_type="bh2", lower_order_final=True,
                 disable_corrector=(), final_sigmas_type="zero", **_ignored):
        if solver_order not in (1, 2) or not predict_x0 or solver_type != "bh2" or prediction_type != "flow_prediction" \
                or use_dynamic_shifting or final_sigmas_type != "zero" or not lower_order_final:
            raise _lib.VCError("FlowUniPCMultistepScheduler: only the configuration the reference uses is implemented "
                               "(flow_prediction, predict_x0, bh2, solver_order <= 2, lower_order_final, final sigma 0)")
        self.num_train_timesteps, self.solver_order, self.shift = num_train_timesteps, solver_order, float(shift)
        self.disable_corrector = list(disable_corrector)
        sig = 1.0 - np.linspace(1, 1 / num_train_timesteps, num_train_timesteps)[::-1]
        sig = (self.shift * sig / (1 + (self.shift - 1) * sig)).astype(np.float32)
        self.sigma_min, self.sigma_max = float(sig[-1]), float(sig[0])
        self.config = dict(num_train_timesteps=num_train_timesteps, solver_order=solver_order, shift=shift)
        self.timesteps = self.sigmas = None
        self._state = None

    def set_timesteps(self, num_inference_steps, device=None, sigmas=None, mu=None, shift=None):
        shift = self.shift if shift is None else float(shift)
        if sigmas is None:
            sigmas = np.linspace(self.sigma_max, self.sigma_min, num_inference_steps + 1)[:-1]
        s = shift * np.asarray(sigmas, dtype=np.float64) / (1 + (shift - 1) * np.asarray(sigmas, dtype=np.float64))
        self.timesteps = torch.from_numpy(s * self.num_train_timesteps).to(device=device, dtype=torch.int64)
        self.sigmas = torch.from_numpy(np.concatenate([s, [0.0]]).astype(np.float32))
        self.num_inference_steps = len(s)
        self._sig = [float(v) for v in self.sigmas]        # the fp32-rounded values, as host floats
        self._step_index, self._lower_order_nums, self._this_order = 0, 0, 1
        self._state = None

    def scale_model_input(self, sample, *a, **k):
        return sample

    @property
    def step_index(self):
        return self._step_index

    # -- per-step scalars ---------------------------------------------------------------------------------------------
    def _coefficients(self):
        i, sg = self._step_index, self._sig
        use_corr = i > 0 and (i - 1) not in self.disable_corrector and self._state is not None and self._state["have_last"]
        c = [sg[i], 0.0, 0.0, 0.0, 0.0, 0.0, 1.0, 0.0, 1.0, 0.0, 0.0, 0.0, 1.0]
        if use_corr:
            oc = self._this_order                                     # the order the previous predictor used
            lam_s0 = _lam(sg[i - 1])
            h = _lam(sg[i]) - lam_s0
            rk = (_lam(sg[i - 2]) - lam_s0) / h if oc == 2 else 1.0
            h_phi_1, B_h, rhos = _bh(h, oc, rk)
            alpha_t = 1.0 - sg[i]
            c[1], c[2], c[3], c[4] = float(oc), sg[i] / sg[i - 1], alpha_t * h_phi_1, alpha_t * B_h
            c[5], c[6], c[7] = (rhos[0], rk, rhos[1]) if oc == 2 else (0.0, 1.0, rhos[0])
        this_order = min(self.solver_order, self.num_inference_steps - i)       # lower_order_final
        op = min(this_order, self._lower_order_nums + 1)
        lam_s0 = _lam(sg[i])
        h = _lam(sg[i + 1]) - lam_s0
        h_phi_1, B_h, _ = _bh(h, 1, 1.0)
        alpha_t = 1.0 - sg[i + 1]
        c[8], c[9], c[10], c[11] = float(op), sg[i + 1] / sg[i], alpha_t * h_phi_1, alpha_t * B_h
        if op == 2:
            c[12] = (_lam(sg[i - 1]) - lam_s0) / h
        return c, op

    def _buffers(self, sample):
        n = sample.numel()
        st = self._state
        if st is None or st["n"] != n or st["bufs"][0].device != sample.device:
            st = dict(n=n, bufs=[torch.empty(n, dtype=torch.float32, device=sample.device) for _ in range(5)],
                      m0=None, m1=None, last=None, have_last=False)
            st["free"] = list(st["bufs"])
            self._state = st
        return st

    def step_cfg(self, pred_uncond, pred_text, sample, guidance_scale=1.0):
        """CFG combine (pred_text=None: none) + corrector + predictor in one kernel -> next sample (bf16, new tensor)."""
        if self.sigmas is None:
            raise _lib.VCError("FlowUniPCMultistepScheduler.step before set_timesteps")
        if self._step_index >= self.num_inference_steps:
            raise _lib.VCError("FlowUniPCMultistepScheduler.step called more often than num_inference_steps")
        tensors = (pred_uncond, sample) + ((pred_text,) if pred_text is not None else ())
        for u in tensors:
            ops._chk(u, "FlowUniPCMultistepScheduler.step input")
            if not u.is_contiguous() or u.numel() != sample.numel():
                raise _lib.VCError("FlowUniPCMultistepScheduler.step: contiguous tensors of one size expected")
        st = self._buffers(sample)
        coef, op = self._coefficients()
        x0_out, last_out = st["free"].pop(), st["free"].pop()
        out = torch.empty_like(sample)
        lib = _lib.load()
        _lib.check(lib.vc_cfg_unipc_step(ops._ptr(pred_uncond), ops._ptr(pred_text), ops._ptr(sample), ops._ptr(st["last"]),
                                         ops._ptr(st["m0"]), ops._ptr(st["m1"]), ops._ptr(out), ops._ptr(x0_out),
                                         ops._ptr(last_out), sample.numel(), float(guidance_scale),
                                         (ctypes.c_float * 13)(*coef), ops._stream()), "vc_cfg_unipc_step")
        ops.launch_count += 1
        for old in (st["m1"], st["last"]):          # rotate: x0 -> m0 -> m1; corrected sample -> last
            if old is not None:
                st["free"].append(old)
        st["m1"], st["m0"], st["last"], st["have_last"] = st["m0"], x0_out, last_out, True
        self._this_order = op
        if self._lower_order_nums < self.solver_order:
            self._lower_order_nums += 1
        self._step_index += 1
        return out

    def step(self, model_output, timestep=None, sample=None, return_dict=True, generator=None, **_ignored):
        prev = self.step_cfg(model_output.contiguous(), None, sample.contiguous())
        return (prev,) if not return_dict else {"prev_sample": prev}


class FlowDPMSolverMultistepScheduler(FlowUniPCMultistepScheduler):
    """The reference's "Flow_DPM++" sampler (inference/versecrafter_inference.py:38,:255;
    `videox_fun.utils.fm_solvers.FlowDPMSolverMultistepScheduler`, un-vendored): data-prediction DPM-Solver++ 2M with the
    midpoint rule, first-order first and last steps.  Its update is algebraically the UniPC `bh2` predictor of the
    same order (expm1(-h) = B(h), (m1 - m0)/r_k = D1), so it runs on the same kernel with the corrector switched off on
    every step."""

    def __init__(self, *a, algorithm_type="dpmsolver++", solver_type="midpoint", **kw):
        if algorithm_type != "dpmsolver++" or solver_type != "midpoint":
            raise _lib.VCError("FlowDPMSolverMultistepScheduler: only dpmsolver++ / midpoint (the reference's setting)")
        super().__init__(*a, **kw)

    def set_timesteps(self, num_inference_steps, device=None, sigmas=None, mu=None, shift=None):
        super().set_timesteps(num_inference_steps, device=device, sigmas=sigmas, mu=mu, shift=shift)
        self.disable_corrector = list(range(self.num_inference_steps))
