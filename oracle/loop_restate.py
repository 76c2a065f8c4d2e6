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
