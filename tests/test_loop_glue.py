"""Per-step glue kernels: the oracle against fixtures produced by the reference's own
geoada_encode_masks / geoada_latent (CPU), the kernels bit-exact against the oracle (GPU), the sampling loop."""
import os

import numpy as np
import pytest
import torch

from oracle import loop_restate as lr

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def pieces():
    return np.load(os.path.join(G, "loop_pieces.npz"))


@pytest.mark.parametrize("name", ["a", "b", "c"])
def test_oracle_matches_reference_mask_encoding(pieces, name):
    mask = torch.from_numpy(pieces[f"{name}_mask"])
    assert torch.equal(lr.geoada_encode_masks(mask), torch.from_numpy(pieces[f"{name}_enc"]))
    assert torch.equal(lr.geoada_encode_masks(mask, n_ref=2), torch.from_numpy(pieces[f"{name}_enc_ref2"]))
    lat = list(torch.from_numpy(pieces[f"{name}_lat"]))
    assert torch.equal(lr.geoada_context(lat, mask), torch.from_numpy(pieces[f"{name}_ctx"]))


def test_source_frame_table_is_torchs_own():
    from versecrafter_b200 import loop
    for n_in, n_out in ((81, 21), (9, 3), (13, 4), (49, 13), (2, 41), (121, 31)):
        idx = loop.nearest_exact_source_frames(n_in, n_out)
        assert idx == lr.nearest_exact_source_frames(n_in, n_out).tolist() and len(idx) == n_out
        assert all(0 <= v < n_in for v in idx) and idx == sorted(idx)


def test_sigma_schedule_host_matches_oracle():
    from versecrafter_b200 import loop
    for n, shift in ((50, 5.0), (30, 16.0), (8, 1.0)):
        s = loop.flow_shift_sigmas(n, shift)
        assert torch.equal(s, lr.flow_shift_sigmas(n, shift)) and len(s) == n + 1
        assert float(s[0]) == 1.0 and float(s[-1]) == 0.0 and bool((s[:-1] > s[1:]).all())


def test_cpu_tensors_are_refused():
    from versecrafter_b200 import _lib, loop
    x = torch.zeros(1, 16, 1, 4, 4, dtype=torch.bfloat16)
    with pytest.raises(_lib.VCError):
        loop.cfg_euler_step(x, x, x, 5.0, -0.1)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["a", "b", "c"])
@pytest.mark.parametrize("n_ref", [0, 2])
def test_control_latents_kernel_bit_exact(pieces, name, n_ref):
    from versecrafter_b200 import loop
    mask = torch.from_numpy(pieces[f"{name}_mask"]).cuda().bfloat16()
    enc = lr.geoada_encode_masks(mask, n_ref=n_ref)
    g = torch.Generator(device="cuda").manual_seed(3)
    lat = [torch.randn(enc.shape[1:], generator=g, device="cuda").expand(16, *enc.shape[1:]).contiguous().bfloat16() * (i + 1)
           for i in range(4)]
    got = loop.geoada_context(lat, mask, n_ref=n_ref)
    want = lr.geoada_context(lat, mask, n_ref=n_ref)
    assert got.shape == want.shape == (128,) + tuple(enc.shape[1:]) and torch.equal(got, want)
    if n_ref == 0:   # and against the reference-generated fixture itself (mask part; values are bf16-exact 0/u patterns)
        ref = torch.from_numpy(pieces[f"{name}_enc"]).cuda()
        assert torch.equal(got[64:].float(), ref.bfloat16().float())


@pytest.mark.gpu
def test_control_latents_full_size():
    """BASELINE c3 shapes: 81 x 480 x 832 mask, 4 control latents [16, 21, 60, 104]."""
    from versecrafter_b200 import loop
    g = torch.Generator(device="cuda").manual_seed(9)
    mask = (torch.rand(1, 81, 480, 832, generator=g, device="cuda") > 0.5).bfloat16()
    lat = [torch.randn(16, 21, 60, 104, generator=g, device="cuda").bfloat16() for _ in range(4)]
    got = loop.geoada_context(lat, mask)
    assert torch.equal(got, lr.geoada_context(lat, mask))


@pytest.mark.gpu
@pytest.mark.parametrize("guidance,dt", [(5.0, -0.0371), (6.0, -0.2), (1.0, -0.01), (3.5, 0.125)])
def test_cfg_euler_step_bit_exact(guidance, dt):
    from versecrafter_b200 import loop
    g = torch.Generator(device="cuda").manual_seed(17)
    shape = (1, 16, 5, 30, 52)
    u, c, x = (torch.randn(shape, generator=g, device="cuda").bfloat16() for _ in range(3))
    guided = torch.empty_like(x)
    got = loop.cfg_euler_step(u, c, x, guidance, dt, guided_out=guided)
    want_g = lr.cfg_combine(u, c, guidance)
    assert torch.equal(guided, want_g)
    assert torch.equal(got, lr.euler_step(want_g, x, 0.5, 0.5 + dt))
    got1 = loop.cfg_euler_step(u, None, x, guidance, dt)          # no CFG: the single prediction is the model output
    assert torch.equal(got1, lr.euler_step(u, x, 0.5, 0.5 + dt))


@pytest.mark.gpu
def test_denoise_loop_matches_stepwise_oracle():
    """3 Euler steps with CFG through the drop-in transformer == the same loop written with the oracle's glue around
    the same transformer calls (the transformer's own parity is tests/test_model_gpu.py)."""
    from oracle import restate
    from versecrafter_b200 import loop
    from versecrafter_b200.transformer import VerseCrafterWanTransformer3DModel
    cfg = dict(dim=256, ffn_dim=512, num_heads=2, num_layers=2, text_dim=64, geoada_in_dim=128)
    sd = restate.synth_state_dict(cfg, seed=4, device="cuda", dtype=torch.bfloat16)
    model = VerseCrafterWanTransformer3DModel(**cfg, device="cuda", dtype=torch.bfloat16, init=False)
    model.load_state_dict(sd, strict=True)
    inp = restate.synth_inputs(cfg, (2, 8, 16), seed=4, device="cuda")
    x0 = inp["x"][:1].bfloat16()
    geo = inp["geoada_context"][0].bfloat16()
    sig = loop.flow_shift_sigmas(3, 5.0)
    seen = []
    out = loop.denoise(model, x0, geo, inp["context"], sig, guidance_scale=5.0, callback=lambda i, t, x: seen.append(float(t[0])))
    assert seen == pytest.approx([float(s) * 1000 for s in sig[:-1]], rel=1e-6)
    assert model.num_inference_steps == 3 and model.current_steps == 2
    x = x0
    for i in range(3):
        t = torch.full((2,), float(sig[i]) * 1000, device="cuda")
        pred = model(x=torch.cat([x, x]), t=t, geoada_context=[geo, geo], context=inp["context"], seq_len=inp["seq_len"])
        x = lr.euler_step(lr.cfg_combine(pred[0:1], pred[1:2], 5.0), x, sig[i], sig[i + 1])
    assert torch.equal(out, x)
