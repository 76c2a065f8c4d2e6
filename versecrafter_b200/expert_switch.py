"""Wan2.2-A14B high/low-noise expert switch behind ONE drop-in transformer object (BASELINE c5).

The reference ships only the YAML for this model family (/root/reference/config/wan2.2/wan_civitai_t2v.yaml:3-7:
`transformer_low_noise_model_subpath`, `transformer_high_noise_model_subpath`, `transformer_combination_type: "moe"`,
`boundary: 0.875`); no code in the reference tree consumes it and `WanVerseCrafterPipeline` takes a single
`transformer` (versecrafter/pipeline/pipeline_wan_versecrafter.py:187-209).  The rule implemented here is the published
Wan2.2 one: a denoising step whose timestep t satisfies  t >= boundary * num_train_timesteps  runs the HIGH-noise
expert, every later (smaller-t) step the LOW-noise expert.  Both experts are complete
`VerseCrafterWanTransformer3DModel`s (2 x 43 GB bf16 fit one B200's 180 GB), so the unchanged sampling loop keeps
calling `transformer(x=..., t=..., geoada_context=..., context=..., seq_len=...)` and never sees the switch.
"""
import os

import torch
import torch.nn as nn

from ._lib import VCError
from .transformer import VerseCrafterWanTransformer3DModel


class VerseCrafterWan22ExpertSwitch(nn.Module):
    """Two experts, one forward signature.  `high_noise_model` serves t >= boundary * num_train_timesteps."""

    def __init__(self, high_noise_model, low_noise_model, boundary=0.875, num_train_timesteps=1000):
        super().__init__()
        if not 0.0 <= boundary <= 1.0:
            raise VCError(f"boundary must lie in [0, 1], got {boundary}")
        self.high_noise_model = high_noise_model
        self.low_noise_model = low_noise_model
        self.boundary = float(boundary)
        self.num_train_timesteps = int(num_train_timesteps)
        self.last_expert = None          # "high" | "low": which expert served the latest call (tests, logging)
        self.calls = {"high": 0, "low": 0}

    # -- routing -----------------------------------------------------------------------------------------------------
    def select(self, t):
        """Expert name for timestep(s) `t` (python number, 0-d or [B] tensor; all entries of a CFG batch are the
        same step, pipeline:890).  A device tensor costs one scalar read-back per step, like the reference's
        TeaCache decision (wan_transformer3d.py:217-220)."""
        if torch.is_tensor(t):
            if t.numel() == 0:
                raise VCError("empty timestep tensor")
            tv = float(t.reshape(-1)[0].item()) if t.numel() == 1 else float(t.max().item())
        else:
            tv = float(t)
        return "high" if tv >= self.boundary * self.num_train_timesteps else "low"

    def expert(self, name):
        return self.high_noise_model if name == "high" else self.low_noise_model

    def forward(self, x, t, geoada_context, context, seq_len, geoada_context_scale=1.0, clip_fea=None, y=None,
                cond_flag=True):
        name = self.select(t)
        self.last_expert = name
        self.calls[name] += 1
        return self.expert(name)(x=x, t=t, geoada_context=geoada_context, context=context, seq_len=seq_len,
                                 geoada_context_scale=geoada_context_scale, clip_fea=clip_fea, y=y,
                                 cond_flag=cond_flag)

    # -- the caller-visible surface of the single-transformer class, applied to both experts -----------
    @property
    def config(self):
        return self.high_noise_model.config

    @property
    def dtype(self):
        return self.high_noise_model.dtype

    @property
    def device(self):
        return self.high_noise_model.device

    @property
    def blocks(self):
        return self.high_noise_model.blocks

    @property
    def freqs(self):
        return self.high_noise_model.freqs

    def _both(self):
        return (self.high_noise_model, self.low_noise_model)

    @property
    def num_inference_steps(self):
        return self.high_noise_model.num_inference_steps

    @num_inference_steps.setter
    def num_inference_steps(self, v):      # pipeline:869
        if "high_noise_model" in self._modules:
            for m in self._both():
                m.num_inference_steps = v

    @property
    def current_steps(self):
        return self.high_noise_model.current_steps

    @current_steps.setter
    def current_steps(self, v):            # pipeline:872
        if "high_noise_model" in self._modules:
            for m in self._both():
                m.current_steps = v

    def check_sp_status(self):
        """Blocking sequence-parallel barrier check after a denoise loop (the experts share one exchange)."""
        self.low_noise_model.check_sp_status()

    def enable_teacache(self, *a, **kw):
        for m in self._both():
            m.enable_teacache(*a, **kw)

    def disable_teacache(self):
        for m in self._both():
            m.disable_teacache()

    def enable_cfg_skip(self, *a, **kw):
        for m in self._both():
            m.enable_cfg_skip(*a, **kw)

    def disable_cfg_skip(self):
        for m in self._both():
            m.disable_cfg_skip()

    def enable_riflex(self, *a, **kw):
        for m in self._both():
            m.enable_riflex(*a, **kw)

    def disable_riflex(self):
        for m in self._both():
            m.disable_riflex()

    def enable_multi_gpus_inference(self, group=None):
        for m in self._both():
            m.enable_multi_gpus_inference(group)
        # one symmetric-memory exchange (receive buffers + barrier pads) serves both experts: only one runs at a time
        # (object.__setattr__: a plain reference, not a registered sub-module of the low-noise expert)
        object.__setattr__(self.low_noise_model, "_px_owner", self.high_noise_model)

    def release_workspace(self):
        for m in self._both():
            m.release_workspace()

    def share_workspace(self):
        """Both experts run strictly one after the other, so they can use one activation workspace (8 GB at c3)."""
        self.low_noise_model._ws = self.high_noise_model._ws
        return self

    # -- loading: <root>/{high_noise_model, low_noise_model}/ each in the single-transformer layout ------------------------
    @classmethod
    def from_pretrained(cls, pretrained_model_path, subfolder=None, transformer_additional_kwargs={},
                        low_cpu_mem_usage=False, torch_dtype=torch.bfloat16):
        kw = dict(transformer_additional_kwargs)
        low = kw.pop("transformer_low_noise_model_subpath", "./low_noise_model")
        high = kw.pop("transformer_high_noise_model_subpath", "./high_noise_model")
        comb = kw.pop("transformer_combination_type", "moe")
        boundary = kw.pop("boundary", 0.875)
        if comb != "moe":
            raise VCError(f"transformer_combination_type {comb!r}: only 'moe' (timestep-boundary switch) exists")
        root = pretrained_model_path if subfolder is None else os.path.join(pretrained_model_path, subfolder)
        models = [VerseCrafterWanTransformer3DModel.from_pretrained(
            os.path.normpath(os.path.join(root, sub)), transformer_additional_kwargs=kw,
            low_cpu_mem_usage=low_cpu_mem_usage, torch_dtype=torch_dtype) for sub in (high, low)]
        return cls(models[0], models[1], boundary=boundary)
