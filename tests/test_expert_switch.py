"""Wan2.2 high/low-noise expert switch: routing rule and caller surface on CPU (stand-in experts),
parity of each routed call against the oracle run with that expert's state dict on the GPU."""
import json
import os

import pytest
import torch
import torch.nn as nn

from versecrafter_b200._lib import VCError
from versecrafter_b200.expert_switch import VerseCrafterWan22ExpertSwitch
from versecrafter_b200.transformer import VerseCrafterWanTransformer3DModel

TINY = dict(dim=128, ffn_dim=256, num_heads=1, num_layers=2, text_dim=32, geoada_in_dim=128)


class _Stub(nn.Module):
    """Records calls; returns a constant so the routed expert is visible in the output."""

    def __init__(self, value):
        super().__init__()
        self.value, self.seen = value, []
        self.num_inference_steps, self.current_steps = None, 0

    def forward(self, x, t, geoada_context, context, seq_len, geoada_context_scale=1.0, clip_fea=None, y=None,
                cond_flag=True):
        self.seen.append((float(t.reshape(-1)[0]), seq_len, geoada_context_scale, cond_flag))
        return torch.full_like(x, self.value)


def test_routing_follows_the_timestep_boundary():
    hi, lo = _Stub(1.0), _Stub(-1.0)
    sw = VerseCrafterWan22ExpertSwitch(hi, lo, boundary=0.875)
    x = torch.zeros(2, 16, 1, 4, 4)
    schedule = [999.0, 900.0, 875.0, 874.9, 500.0, 20.0]
    outs = [sw(x=x, t=torch.full((2,), tv), geoada_context=x, context=[x, x], seq_len=4, geoada_context_scale=0.7)
            for tv in schedule]
    assert [float(o[0, 0, 0, 0, 0]) for o in outs] == [1.0, 1.0, 1.0, -1.0, -1.0, -1.0]
    assert sw.calls == {"high": 3, "low": 3} and sw.last_expert == "low"
    assert [s[0] for s in hi.seen] == [999.0, 900.0, 875.0] and hi.seen[0][1:] == (4, 0.7, True)
    assert sw.select(875) == "high" and sw.select(torch.tensor(874.0)) == "low"
    with pytest.raises(VCError):
        VerseCrafterWan22ExpertSwitch(hi, lo, boundary=1.5)


def test_pipeline_attributes_reach_both_experts():
    hi = VerseCrafterWanTransformer3DModel(**TINY)
    lo = VerseCrafterWanTransformer3DModel(**TINY)
    sw = VerseCrafterWan22ExpertSwitch(hi, lo)
    sw.num_inference_steps = 50          # pipeline:869
    sw.current_steps = 7                 # pipeline:872
    assert (hi.num_inference_steps, lo.num_inference_steps, hi.current_steps, lo.current_steps) == (50, 50, 7, 7)
    assert sw.config.patch_size == (1, 2, 2) and len(sw.blocks) == 2 and sw.freqs.shape == (1024, 64)
    sw.enable_teacache([1.0, 0.0], 10, 0.1, num_skip_start_steps=2, offload=False)
    sw.enable_cfg_skip(0.25, 40)
    assert hi.teacache is not None and lo.teacache is not None and hi.teacache is not lo.teacache
    assert lo.cfg_skip_ratio == 0.25
    sw.disable_teacache(); sw.disable_cfg_skip()
    assert hi.teacache is None and lo.cfg_skip_ratio is None
    f0 = lo.freqs.clone()
    sw.enable_riflex(k=6, L_test=30)
    assert not torch.equal(lo.freqs, f0)
    sw.disable_riflex()
    assert torch.equal(lo.freqs, f0)
    # state dict: both experts under their own prefixes, nothing shared or duplicated
    keys = set(sw.state_dict())
    assert {k.split(".", 1)[0] for k in keys} == {"high_noise_model", "low_noise_model"}
    assert len(keys) == 2 * len(hi.state_dict())
    sw.share_workspace()
    assert lo._ws is hi._ws


def test_from_pretrained_reads_the_wan22_layout(tmp_path):
    """<root>/high_noise_model and <root>/low_noise_model, each a single-transformer folder
    (config/wan2.2/wan_civitai_t2v.yaml:3-7)."""
    from safetensors.torch import save_file
    sds = {}
    for sub, seed in (("high_noise_model", 1), ("low_noise_model", 2)):
        torch.manual_seed(seed)
        m = VerseCrafterWanTransformer3DModel(**TINY)
        d = tmp_path / sub
        d.mkdir()
        with open(d / "config.json", "w") as f:
            json.dump(TINY, f)
        sds[sub] = {k: v.clone().contiguous() for k, v in m.state_dict().items()}
        save_file(sds[sub], str(d / "diffusion_pytorch_model.safetensors"))
    kw = dict(transformer_low_noise_model_subpath="./low_noise_model",
              transformer_high_noise_model_subpath="./high_noise_model", transformer_combination_type="moe",
              boundary=0.9)
    sw = VerseCrafterWan22ExpertSwitch.from_pretrained(str(tmp_path), transformer_additional_kwargs=kw,
                                                       torch_dtype=torch.float32)
    assert sw.boundary == 0.9
    for sub in sds:
        got = getattr(sw, sub).state_dict()
        assert all(torch.equal(got[k], v) for k, v in sds[sub].items())
    with pytest.raises(VCError):
        VerseCrafterWan22ExpertSwitch.from_pretrained(str(tmp_path), transformer_additional_kwargs=dict(
            kw, transformer_combination_type="concat"))


@pytest.mark.gpu
def test_each_step_matches_the_oracle_of_the_routed_expert():
    from oracle import restate
    cfg = dict(dim=256, ffn_dim=512, num_heads=2, num_layers=2, text_dim=64, geoada_in_dim=128)
    experts, sds = [], []
    for seed in (21, 22):
        sd = restate.synth_state_dict(cfg, seed=seed, device="cuda", dtype=torch.bfloat16)
        m = VerseCrafterWanTransformer3DModel(**cfg, device="cuda", dtype=torch.bfloat16, init=False)
        m.load_state_dict(sd, strict=True)
        experts.append(m.eval()); sds.append(sd)
    sw = VerseCrafterWan22ExpertSwitch(experts[0], experts[1], boundary=0.875).share_workspace()
    for t_value, which in ((950.0, 0), (875.0, 0), (600.0, 1), (30.0, 1)):
        inp = restate.synth_inputs(cfg, (2, 6, 10), seed=5, device="cuda", t_value=t_value)
        y = sw(x=inp["x"].bfloat16(), t=inp["t"], geoada_context=inp["geoada_context"].bfloat16(),
               context=inp["context"], seq_len=inp["seq_len"])
        assert sw.last_expert == ("high", "low")[which]
        ref = {}
        for i in (0, 1):
            ref[i] = restate.forward({k: v.float() for k, v in sds[i].items()}, cfg, inp["x"], inp["t"],
                                     inp["geoada_context"], inp["context"], inp["seq_len"], dtype=torch.float32)
        cos = lambda a, b: float(torch.nn.functional.cosine_similarity(a.double().flatten(), b.double().flatten(), dim=0))
        assert cos(y, ref[which]) >= 0.999
        assert cos(y, ref[1 - which]) < 0.99   # the other expert's weights give a different prediction
