"""TEST INFRASTRUCTURE — CPU/torch restatement of the per-step glue either side of the transformer forward
(SURVEY.md §8f rows N1, N3).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline may import this.

Pinning: geoada_encode_masks / geoada_latent are pinned to the reference's own functions (extracted from
/root/reference/versecrafter/pipeline/pipeline_wan_versecrafter.py by tests/golden/gen_golden_loop.py ->
tests/golden/loop_pieces.npz).  The CFG combine is the pipeline's own expression (:905-906).  The Euler update follows
diffusers' FlowMatchEulerDiscreteScheduler.step, which is NOT in the reference tree and not installed here:
**parity unpinned** for that function (restated from the published algorithm)."""
import torch
import torch.nn.functional as F


def geoada_encode_masks(mask, n_ref=0, vae_stride=(4, 8, 8)):
    """mask [c, T, H, W] -> [64, n_ref + (T+3)//4, H/8, W/8]   (pipeline_wan_versecrafter.py:440-486)."""
    c, depth, height, width = mask.shape
    new_depth = int((depth + 3) // vae_stride[0])
    height = 2 * (int(height) // (vae_stride[1] * 2))
    width = 2 * (int(width) // (vae_stride[2] * 2))
    m = mask[0].view(depth, height, vae_stride[1], width, vae_stride[1]).permute(2, 4, 0, 1, 3)
    m = m.reshape(vae_stride[1] * vae_stride[2], depth, height, width)
    m = F.interpolate(m.unsqueeze(0).float(), size=(new_depth, height, width), mode="nearest-exact").squeeze(0).to(mask.dtype)
    if n_ref:
        m = torch.cat((torch.zeros_like(m[:, :n_ref]), m), dim=1)
    return m


def geoada_context(latents, mask, n_ref=0):
    """latents: list of [16, F, H', W'] VAE latents of the control videos (channel-concatenated, :420-424) ->
    [len*16 + 64, F, H', W']  (geoada_latent, :488-489)."""
    return torch.cat([torch.cat(list(latents), dim=0), geoada_encode_masks(mask, n_ref)], dim=0)


def nearest_exact_source_frames(n_in, n_out):
    """Source index of every output frame of F.interpolate(mode='nearest-exact') along one axis."""
    x = torch.arange(n_in, dtype=torch.float32).view(1, 1, n_in, 1, 1)
    return F.interpolate(x, size=(n_out, 1, 1), mode="nearest-exact").flatten().long()


def cfg_combine(pred_uncond, pred_text, guidance_scale):
    """pipeline_wan_versecrafter.py:905-906 — tensor arithmetic in the tensors' own dtype (bf16: one rounding per op)."""
    return pred_uncond + guidance_scale * (pred_text - pred_uncond)


def euler_step(model_output, sample, sigma, sigma_next):
    """FlowMatchEulerDiscreteScheduler.step (diffusers; parity unpinned): sample.float() + dt * model_output, cast
    back to model_output.dtype; dt is a 0-dim fp32 tensor, so dt * model_output keeps model_output's dtype."""
    dt = torch.tensor(float(sigma_next) - float(sigma), dtype=torch.float32, device=sample.device)
    prev = sample.to(torch.float32) + dt * model_output
    return prev.to(model_output.dtype)


def flow_shift_sigmas(num_steps, shift=5.0, num_train_timesteps=1000):
    """Shifted flow-matching sigma schedule (FlowMatchEulerDiscreteScheduler.set_timesteps without dynamic shifting;
    parity unpinned): sigma in linspace(1, 1/N_train, steps) -> shift*s / (1 + (shift-1)*s), then a final 0."""
    def sh(s):
        return shift * s / (1 + (shift - 1) * s)
    smax, smin = sh(torch.tensor(1.0, dtype=torch.float64)), sh(torch.tensor(1.0 / num_train_timesteps, dtype=torch.float64))
    t = torch.linspace(float(smax) * num_train_timesteps, float(smin) * num_train_timesteps, num_steps, dtype=torch.float64)
    s = sh(t / num_train_timesteps)
    return torch.cat([s, torch.zeros(1, dtype=torch.float64)]).float()


class FlowUniPCOracle:
    """Flow-matching UniPC multistep scheduler (predict-x0, `bh2`, solver order 2, lower_order_final, final sigma 0) —
    the reference's default sampler (`sampler_name = "Flow_Unipc"`, /root/reference/inference/versecrafter_inference.py
    :135,:254; driven at pipeline_wan_versecrafter.py:750-752,:909).  The class itself is
    `videox_fun.utils.fm_solvers_unipc.FlowUniPCMultistepScheduler` (VideoX-Fun, un-vendored, no recoverable pin; a
    flow-matching adaptation of diffusers' UniPCMultistepScheduler): **parity unpinned**.  This restates the published
    UniPC algorithm (Zhao et al. 2023, B(h) = expm1(h) variant) in the structure of that scheduler — model-output
    history list, per-step R/b system, einsum over the D1 stack — in fp32 torch ops."""

    def __init__(self, num_train_timesteps=1000, solver_order=2, shift=1.0):
        self.N, self.solver_order, self.shift = num_train_timesteps, solver_order, shift
        alphas = torch.linspace(1, 1 / num_train_timesteps, num_train_timesteps, dtype=torch.float64).flip(0)
        sig = 1.0 - alphas
        sig = shift * sig / (1 + (shift - 1) * sig)
        self.sigma_min, self.sigma_max = float(sig[-1].float()), float(sig[0].float())

    def set_timesteps(self, num_inference_steps, shift=None):
        shift = self.shift if shift is None else shift
        import numpy as np
        s = np.linspace(self.sigma_max, self.sigma_min, num_inference_steps + 1)[:-1]
        s = shift * s / (1 + (shift - 1) * s)
        self.timesteps = torch.from_numpy(s * self.N).to(torch.int64)
        self.sigmas = torch.from_numpy(np.concatenate([s, [0.0]]).astype(np.float32))
        self.model_outputs = [None] * self.solver_order
        self.lower_order_nums, self.last_sample, self.step_index, self.this_order = 0, None, 0, 1

    @staticmethod
    def _lam(sigma):
        return torch.log(1 - sigma) - torch.log(sigma)

    def _rb(self, rks, hh, order):
        h_phi_1 = torch.expm1(hh)
        h_phi_k = h_phi_1 / hh - 1
        B_h = torch.expm1(hh)
        R, b, fact = [], [], 1
        for i in range(1, order + 1):
            R.append(torch.pow(rks, i - 1))
            b.append(h_phi_k * fact / B_h)
            fact *= i + 1
            h_phi_k = h_phi_k / hh - 1 / fact
        return torch.stack(R), torch.stack(b), h_phi_1, B_h

    def _d1s(self, first_back, order, lam_s0, h, m0):
        rks, d1s = [], []
        for i in range(1, order):
            si = self.step_index - first_back - (i - 1)
            mi = self.model_outputs[-(i + 1)]
            rk = (self._lam(self.sigmas[si]) - lam_s0) / h
            rks.append(rk)
            d1s.append((mi - m0) / rk)
        rks.append(torch.tensor(1.0))
        return torch.stack(rks), (torch.stack(d1s, dim=1) if d1s else None)

    def _predict(self, sample, order):
        m0, x = self.model_outputs[-1], sample
        sigma_t, sigma_s0 = self.sigmas[self.step_index + 1], self.sigmas[self.step_index]
        alpha_t = 1 - sigma_t
        lam_s0 = self._lam(sigma_s0)
        h = self._lam(sigma_t) - lam_s0
        rks, D1s = self._d1s(1, order, lam_s0, h, m0)
        R, b, h_phi_1, B_h = self._rb(rks, -h, order)
        x_t_ = sigma_t / sigma_s0 * x - alpha_t * h_phi_1 * m0
        if D1s is None:
            return x_t_
        rhos_p = torch.tensor([0.5]) if order == 2 else torch.linalg.solve(R[:-1, :-1], b[:-1])
        pred_res = torch.einsum("k,bkc...->bc...", rhos_p, D1s)
        return x_t_ - alpha_t * B_h * pred_res

    def _correct(self, model_t, last_sample, order):
        m0, x = self.model_outputs[-1], last_sample
        sigma_t, sigma_s0 = self.sigmas[self.step_index], self.sigmas[self.step_index - 1]
        alpha_t = 1 - sigma_t
        lam_s0 = self._lam(sigma_s0)
        h = self._lam(sigma_t) - lam_s0
        rks, D1s = self._d1s(2, order, lam_s0, h, m0)
        R, b, h_phi_1, B_h = self._rb(rks, -h, order)
        rhos_c = torch.tensor([0.5]) if order == 1 else torch.linalg.solve(R, b)
        x_t_ = sigma_t / sigma_s0 * x - alpha_t * h_phi_1 * m0
        corr_res = torch.einsum("k,bkc...->bc...", rhos_c[:-1], D1s) if D1s is not None else 0
        return x_t_ - alpha_t * B_h * (corr_res + rhos_c[-1] * (model_t - m0))

    def step(self, model_output, sample):
        """model_output = the (CFG-combined) velocity prediction at sigmas[step_index]; returns the next sample in
        sample.dtype.  State (x0 history, last sample) is fp32."""
        out_dtype = sample.dtype
        sample = sample.float()
        x0 = sample - self.sigmas[self.step_index] * model_output.float()           # flow prediction -> x0
        if self.step_index > 0 and self.last_sample is not None:
            sample = self._correct(x0, self.last_sample, self.this_order)
        self.model_outputs = self.model_outputs[1:] + [x0]
        this_order = min(self.solver_order, len(self.timesteps) - self.step_index)   # lower_order_final
        self.this_order = min(this_order, self.lower_order_nums + 1)
        self.last_sample = sample
        prev = self._predict(sample, self.this_order)
        if self.lower_order_nums < self.solver_order:
            self.lower_order_nums += 1
        self.step_index += 1
        return prev.to(out_dtype)


class FlowDPMSolverOracle:
    """Flow-matching DPM-Solver++ multistep scheduler (2M, midpoint, lower_order_final, final sigma 0) — the reference's
    third sampler option "Flow_DPM++" (/root/reference/inference/versecrafter_inference.py:38,:255;
    `videox_fun.utils.fm_solvers.FlowDPMSolverMultistepScheduler`, un-vendored: **parity unpinned**).  Restates the
    published DPM-Solver++ update (Lu et al. 2022, data-prediction form) in the structure of diffusers'
    DPMSolverMultistepScheduler (first-order update / multistep second-order update), fp32 torch ops; shares the sigma
    schedule with FlowUniPCOracle."""

    def __init__(self, num_train_timesteps=1000, solver_order=2, shift=1.0):
        self._sched = FlowUniPCOracle(num_train_timesteps, solver_order, shift)
        self.solver_order = solver_order

    def set_timesteps(self, num_inference_steps, shift=None):
        self._sched.set_timesteps(num_inference_steps, shift)
        self.timesteps, self.sigmas = self._sched.timesteps, self._sched.sigmas
        self.model_outputs = [None] * self.solver_order
        self.lower_order_nums, self.step_index = 0, 0

    @staticmethod
    def _lam(sigma):
        return torch.log(1 - sigma) - torch.log(sigma)

    def _first_order(self, sample):
        sigma_t, sigma_s = self.sigmas[self.step_index + 1], self.sigmas[self.step_index]
        h = self._lam(sigma_t) - self._lam(sigma_s)
        return (sigma_t / sigma_s) * sample - ((1 - sigma_t) * (torch.exp(-h) - 1.0)) * self.model_outputs[-1]

    def _second_order(self, sample):
        sigma_t, sigma_s0, sigma_s1 = (self.sigmas[self.step_index + 1], self.sigmas[self.step_index],
                                       self.sigmas[self.step_index - 1])
        lam_t, lam_s0, lam_s1 = self._lam(sigma_t), self._lam(sigma_s0), self._lam(sigma_s1)
        m0, m1 = self.model_outputs[-1], self.model_outputs[-2]
        h, h_0 = lam_t - lam_s0, lam_s0 - lam_s1
        r0 = h_0 / h
        D0, D1 = m0, (1.0 / r0) * (m0 - m1)
        alpha_t = 1 - sigma_t
        return (sigma_t / sigma_s0) * sample - (alpha_t * (torch.exp(-h) - 1.0)) * D0 \
            - 0.5 * (alpha_t * (torch.exp(-h) - 1.0)) * D1

    def step(self, model_output, sample):
        out_dtype = sample.dtype
        sample = sample.float()
        x0 = sample - self.sigmas[self.step_index] * model_output.float()
        self.model_outputs = self.model_outputs[1:] + [x0]
        final = self.step_index == len(self.timesteps) - 1           # final sigma is 0: first-order last step
        if self.solver_order == 1 or self.lower_order_nums < 1 or final:
            prev = self._first_order(sample)
        else:
            prev = self._second_order(sample)
        if self.lower_order_nums < self.solver_order:
            self.lower_order_nums += 1
        self.step_index += 1
        return prev.to(out_dtype)
