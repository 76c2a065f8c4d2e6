"""The per-step glue either side of the transformer forward, as kernels of libvc_b200.so:

  * geoada_context(): the GeoAdapter input [4*16 + 64, F, H', W'] from the control videos' VAE latents and the merged
    mask — `geoada_encode_multi_frames`' channel concat, `geoada_encode_masks`, `geoada_latent`
    (/root/reference/versecrafter/pipeline/pipeline_wan_versecrafter.py:420-424, :440-486, :488-489) in one pass;
  * cfg_euler_step(): CFG combine (:903-906) + flow-matching Euler update (`scheduler.step`, :909) in one pass;
  * denoise(): the sampling loop of `WanVerseCrafterPipeline.__call__` (:866-909) around a drop-in transformer
    (single model or the Wan2.2 expert switch) for callers that do not bring the reference pipeline object.

The reference's default sampler (flow-matching UniPC) is schedulers.FlowUniPCMultistepScheduler (one fused kernel per
step), schedulers.FlowDPMSolverMultistepScheduler the "Flow_DPM++" one; pass either to denoise(scheduler=...).
The VAE and the text encoder stay external."""
import ctypes

import torch
import torch.nn.functional as F

from . import _lib, ops


def nearest_exact_source_frames(n_in, n_out):
    """Source index per output frame of F.interpolate(mode='nearest-exact') — taken from torch's own routine so the
    kernel's resampling is the reference's by construction (:470)."""
    x = torch.arange(n_in, dtype=torch.float32).view(1, 1, n_in, 1, 1)
    return [int(v) for v in F.interpolate(x, size=(n_out, 1, 1), mode="nearest-exact").flatten().tolist()]


def geoada_context(latents, mask, n_ref=0, vae_stride=(4, 8, 8), out=None):
    """latents: list of bf16 CUDA tensors [16, F, H', W'] (VAE latents of the control videos, reference-image frames
    first when n_ref > 0); mask: bf16 CUDA [c, T, 8H', 8W'] (channel 0 is used, as in the reference) ->
    bf16 [len(latents)*16 + 64, F, H', W'] with F = n_ref + (T + 3) // 4."""
    if tuple(vae_stride[1:]) != (8, 8):
        raise _lib.VCError("geoada_context: only the Wan VAE's 8x8 spatial stride is implemented")
    lat = [u.contiguous() for u in latents]
    for u in lat + [mask]:
        ops._chk(u, "geoada_context input")
    c_lat, Fl, Hl, Wl = lat[0].shape
    _, T, H, W = mask.shape
    if any(u.shape != lat[0].shape for u in lat):
        raise _lib.VCError("geoada_context: all control latents must share one shape")
    if (H, W) != (8 * Hl, 8 * Wl) or H % 16 or W % 16:   # height = 2 * (H // 16) in the reference (:459-460)
        raise _lib.VCError(f"geoada_context: mask {H}x{W} does not match latents {Hl}x{Wl} x 8 (multiples of 16)")
    new_depth = (T + 3) // vae_stride[0]
    if Fl != n_ref + new_depth:
        raise _lib.VCError(f"geoada_context: latents have {Fl} frames, mask resamples to {n_ref} + {new_depth}")
    plane = mask[0].contiguous()
    idx = nearest_exact_source_frames(T, new_depth)
    if out is None:
        out = torch.empty((len(lat) * c_lat + 64, Fl, Hl, Wl), dtype=torch.bfloat16, device=mask.device)
    ptrs = (ctypes.c_void_p * len(lat))(*[u.data_ptr() for u in lat])
    lib = _lib.load()
    _lib.check(lib.vc_control_latents(ptrs, len(lat), c_lat, ops._ptr(plane), T, Fl, Hl, Wl, n_ref,
                                      (ctypes.c_int * len(idx))(*idx), ops._ptr(out), ops._stream()), "vc_control_latents")
    ops.launch_count += 1
    return out


def cfg_euler_step(pred_uncond, pred_text, latents, guidance_scale, dt, out=None, guided_out=None):
    """latents + dt * (u + g (c - u)) with the reference's bf16 rounding points (see include/vc_b200.h);
    pred_text=None -> no CFG."""
    for u in (pred_uncond, latents) + ((pred_text,) if pred_text is not None else ()):
        ops._chk(u, "cfg_euler_step input")
        if not u.is_contiguous() or u.numel() != latents.numel():
            raise _lib.VCError("cfg_euler_step: contiguous tensors of one size expected")
    if out is None:
        out = torch.empty_like(latents)
    # `dt * model_output` in scheduler.step multiplies a 0-dim fp32 tensor with a bf16 tensor: torch's type promotion
    # casts dt to bf16 first, so the step size the reference really applies is bf16(dt)
    dt = float(torch.tensor(float(dt), dtype=torch.float32).to(latents.dtype))
    lib = _lib.load()
    _lib.check(lib.vc_cfg_euler_step(ops._ptr(pred_uncond), ops._ptr(pred_text), ops._ptr(latents), ops._ptr(out),
                                     ops._ptr(guided_out), latents.numel(), float(guidance_scale), float(dt),
                                     ops._stream()), "vc_cfg_euler_step")
    ops.launch_count += 1
    return out


def flow_shift_sigmas(num_steps, shift=5.0, num_train_timesteps=1000):
    """Shifted flow-matching sigmas, N+1 values ending in 0 (FlowMatchEulerDiscreteScheduler.set_timesteps without
    dynamic shifting, diffusers; the reference configures `shift` in config/*/wan_civitai*.yaml scheduler_kwargs)."""
    def sh(s):
        return shift * s / (1 + (shift - 1) * s)
    t = torch.linspace(float(sh(torch.tensor(1.0, dtype=torch.float64))) * num_train_timesteps,
                       float(sh(torch.tensor(1.0 / num_train_timesteps, dtype=torch.float64))) * num_train_timesteps,
                       num_steps, dtype=torch.float64)
    return torch.cat([sh(t / num_train_timesteps), torch.zeros(1, dtype=torch.float64)]).float()


@torch.no_grad()
def denoise(transformer, latents, geoada_ctx, context, sigmas=None, guidance_scale=5.0, geoada_context_scale=1.0,
            num_train_timesteps=1000, callback=None, scheduler=None):
    """The denoising loop of pipeline_wan_versecrafter.py:866-909.  latents [1, 16, F, H, W] bf16; geoada_ctx
    [128, F, H, W]; context = [negative, positive] text embeddings (CFG) or [positive].  Either `sigmas` (N+1
    descending values, flow_shift_sigmas: Euler flow-matching updates) or `scheduler` (a
    schedulers.FlowUniPCMultistepScheduler after set_timesteps: the reference's default sampler; the model sees its
    integer `timesteps`, :750-752,:890).  Returns the final latents."""
    do_cfg = guidance_scale > 1.0 and len(context) == 2
    B = 2 if do_cfg else 1
    _, _, Fr, Hh, Ww = latents.shape
    ps = transformer.config.patch_size
    seq_len = -(-(Hh * Ww) // (ps[1] * ps[2])) * Fr                     # :858-862
    if (sigmas is None) == (scheduler is None):
        raise _lib.VCError("denoise: pass exactly one of `sigmas` (Euler) or `scheduler` (UniPC)")
    n_steps = len(sigmas) - 1 if scheduler is None else len(scheduler.timesteps)
    transformer.num_inference_steps = n_steps                           # :869
    geo = [geoada_ctx] * B                                              # :883-887 (both CFG halves share the control)
    x = latents.contiguous()
    for i in range(n_steps):
        transformer.current_steps = i                                   # :872
        if scheduler is None:
            t = torch.full((B,), float(sigmas[i]) * num_train_timesteps, device=x.device, dtype=torch.float32)
        else:
            t = scheduler.timesteps[i].to(x.device).expand(B)
        pred = transformer(x=torch.cat([x] * B) if do_cfg else x, t=t, geoada_context=geo, context=context,
                           seq_len=seq_len, geoada_context_scale=geoada_context_scale)
        if scheduler is None:
            dt = float(sigmas[i + 1]) - float(sigmas[i])
            x = cfg_euler_step(pred[0:1], pred[1:2] if do_cfg else None, x, guidance_scale, dt)
        else:
            x = scheduler.step_cfg(pred[0:1], pred[1:2] if do_cfg else None, x, guidance_scale)
        if callback is not None:
            callback(i, t, x)
    if hasattr(transformer, "check_sp_status"):
        transformer.check_sp_status()   # sequence parallelism: no barrier of the loop timed out (blocking, once per loop)
    return x
