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


# ---- flow-matching UniPC (the reference's default sampler; upstream class un-vendored: parity unpinned) ----------
def _kernel_formula(coef, u, c, x, last, m0, m1, guidance):
    """The arithmetic of cfg_unipc_step_kernel (csrc/loop_glue.cu) written with torch fp32 ops in the same order."""
    f = lambda v: torch.tensor(v, dtype=torch.float32)
    v = u if c is None else lr.cfg_combine(u, c, guidance)
    x, v = x.float(), v.float()
    x0 = x - f(coef[0]) * v
    xc = x
    if int(coef[1]) >= 1:
        res = f(coef[7]) * (x0 - m0)
        if int(coef[1]) == 2:
            res = f(coef[5]) * ((m1 - m0) / f(coef[6])) + res
        xc = (f(coef[2]) * last - f(coef[3]) * m0) - f(coef[4]) * res
    out = f(coef[9]) * xc - f(coef[10]) * x0
    if int(coef[8]) == 2:
        out = out - f(coef[11]) * (f(0.5) * ((m0 - x0) / f(coef[12])))
    return out, x0, xc


def test_unipc_oracle_properties():
    """An exact straight-path velocity is integrated exactly at every order; a nonlinear field converges at order 2."""
    torch.manual_seed(0)
    x0, n = torch.randn(1, 4, 3, 8, 8), torch.randn(1, 4, 3, 8, 8)
    for steps, shift in ((10, 16.0), (5, 1.0), (3, 5.0), (2, 3.0), (1, 3.0)):
        s = lr.FlowUniPCOracle()
        s.set_timesteps(steps, shift=shift)
        assert s.timesteps[0] == 999 and float(s.sigmas[-1]) == 0.0 and len(s.sigmas) == steps + 1
        x = (1 - s.sigmas[0]) * x0 + s.sigmas[0] * n
        for _ in range(steps):
            x = s.step(n - x0, x)
        assert float((x - x0).abs().max()) < 5e-6

    def run(steps):
        s = lr.FlowUniPCOracle()
        s.set_timesteps(steps, shift=1.0)
        x = torch.ones(1, 1, 1, 1, 1)
        for i in range(steps):
            x = s.step(torch.sin(3 * x) + s.sigmas[i], x)
        return float(x)
    ref = run(2000)
    e = [abs(run(k) - ref) for k in (10, 20, 40)]
    assert 3.0 < e[0] / e[1] < 5.0 and 3.0 < e[1] / e[2] < 5.0        # error / 4 when the step halves


@pytest.mark.parametrize("steps,shift,order", [(8, 16.0, 2), (3, 5.0, 2), (2, 1.0, 2), (1, 3.0, 2), (6, 3.0, 1)])
def test_unipc_host_coefficients_match_oracle(steps, shift, order):
    """schedulers.FlowUniPCMultistepScheduler's per-step scalars, pushed through the kernel's formula on CPU, follow
    the oracle over a whole loop (timesteps / sigmas identical)."""
    from versecrafter_b200 import schedulers
    sch = schedulers.FlowUniPCMultistepScheduler(shift=1, solver_order=order)
    sch.set_timesteps(steps, shift=shift)
    orc = lr.FlowUniPCOracle(solver_order=order)
    orc.set_timesteps(steps, shift=shift)
    assert torch.equal(sch.timesteps, orc.timesteps) and torch.equal(sch.sigmas, orc.sigmas)
    assert sch.order == 1 and sch.init_noise_sigma == 1.0
    g = torch.Generator().manual_seed(steps)
    x = torch.randn(1, 16, 2, 6, 8, generator=g).bfloat16()
    last = m0 = m1 = None
    sch._state = dict(have_last=False)
    for i in range(steps):
        u, c = (torch.randn(x.shape, generator=g).bfloat16() for _ in range(2))
        coef, op = sch._coefficients()
        out, x0, xc = _kernel_formula(coef, u, c, x, last, m0, m1, 5.0)
        m1, m0, last = m0, x0, xc
        sch._state["have_last"] = True
        sch._this_order = op
        sch._lower_order_nums = min(sch._lower_order_nums + 1, sch.solver_order)
        sch._step_index += 1
        want = orc.step(lr.cfg_combine(u, c, 5.0), x)       # same input sample on both sides: per-step agreement
        assert torch.allclose(x0, orc.model_outputs[-1], rtol=1e-5, atol=1e-5)
        assert torch.allclose(xc, orc.last_sample, rtol=1e-4, atol=1e-4)
        x = out.bfloat16()
        d = (x.float() - want.float()).abs()                # fp32 values a few ulps apart -> rare 1-ulp bf16 flips
        assert float(d.max()) <= 2 ** -7 * max(1.0, float(want.float().abs().max())) and float((d > 0).float().mean()) < 0.01


@pytest.mark.gpu
@pytest.mark.parametrize("steps,shift,cfg", [(8, 16.0, True), (3, 5.0, True), (1, 3.0, False), (5, 1.0, False)])
def test_unipc_kernel_matches_oracle(steps, shift, cfg):
    """vc_cfg_unipc_step over a whole loop == the oracle (fp32 state on both sides; host scalars are float64 on one
    side and fp32 torch on the other: agreement to a few fp32 ulps before the bf16 output rounding)."""
    from versecrafter_b200 import schedulers
    sch = schedulers.FlowUniPCMultistepScheduler(shift=1)
    sch.set_timesteps(steps, device="cuda", shift=shift)
    orc = lr.FlowUniPCOracle()
    orc.set_timesteps(steps, shift=shift)
    assert torch.equal(sch.timesteps.cpu(), orc.timesteps)
    g = torch.Generator(device="cuda").manual_seed(31)
    shape = (1, 16, 5, 30, 52)
    x = torch.randn(shape, generator=g, device="cuda").bfloat16()
    for i in range(steps):
        u, c = (torch.randn(shape, generator=g, device="cuda").bfloat16() for _ in range(2))
        got = sch.step_cfg(u, c if cfg else None, x, 5.0)
        v = lr.cfg_combine(u, c, 5.0) if cfg else u
        want = orc.step(v.cpu(), x.cpu())
        assert torch.allclose(sch._state["m0"].view(shape).cpu(), orc.model_outputs[-1], rtol=1e-5, atol=1e-5)
        assert torch.allclose(sch._state["last"].view(shape).cpu(), orc.last_sample, rtol=1e-4, atol=1e-4)
        d = (got.cpu().float() - want.float()).abs()
        assert float(d.max()) <= 2 ** -7 * max(1.0, float(want.float().abs().max()))      # at most one bf16 ulp
        assert float((d > 0).float().mean()) < 0.01
        x = got
    with pytest.raises(Exception):
        sch.step_cfg(u, None, x)                        # more steps than set_timesteps announced
    # the pipeline-compatible entry: step(model_output, t, sample, return_dict=False)
    sch.set_timesteps(2, device="cuda", shift=shift)
    y = sch.step(u, sch.timesteps[0], x, return_dict=False)[0]
    assert y.shape == x.shape and y.dtype == torch.bfloat16 and bool(torch.isfinite(y.float()).all())


@pytest.mark.gpu
def test_denoise_loop_unipc():
    """denoise(scheduler=UniPC) == the same loop written with the oracle scheduler around the same transformer calls."""
    from oracle import restate
    from versecrafter_b200 import loop, schedulers
    from versecrafter_b200.transformer import VerseCrafterWanTransformer3DModel
    cfg = dict(dim=256, ffn_dim=512, num_heads=2, num_layers=2, text_dim=64, geoada_in_dim=128)
    sd = restate.synth_state_dict(cfg, seed=4, device="cuda", dtype=torch.bfloat16)
    model = VerseCrafterWanTransformer3DModel(**cfg, device="cuda", dtype=torch.bfloat16, init=False)
    model.load_state_dict(sd, strict=True)
    inp = restate.synth_inputs(cfg, (2, 8, 16), seed=4, device="cuda")
    x0 = inp["x"][:1].bfloat16()
    geo = inp["geoada_context"][0].bfloat16()
    sch = schedulers.FlowUniPCMultistepScheduler(shift=1)
    sch.set_timesteps(4, device="cuda", shift=16.0)
    out = loop.denoise(model, x0, geo, inp["context"], guidance_scale=5.0, scheduler=sch)
    orc = lr.FlowUniPCOracle()
    orc.set_timesteps(4, shift=16.0)
    x = x0
    for i in range(4):
        t = orc.timesteps[i].cuda().expand(2)
        pred = model(x=torch.cat([x, x]), t=t, geoada_context=[geo, geo], context=inp["context"], seq_len=inp["seq_len"])
        x = orc.step(lr.cfg_combine(pred[0:1], pred[1:2], 5.0).cpu(), x.cpu()).cuda()
    cos = torch.nn.functional.cosine_similarity(out.float().flatten(), x.float().flatten(), dim=0)
    assert float(cos) > 0.9999
    with pytest.raises(Exception):
        loop.denoise(model, x0, geo, inp["context"])     # neither sigmas nor scheduler


# ---- flow-matching DPM-Solver++ 2M ("Flow_DPM++"; upstream class un-vendored: parity unpinned) -------------------
def test_dpm_oracle_integrates_straight_paths_exactly():
    torch.manual_seed(1)
    x0, n = torch.randn(1, 4, 3, 8, 8), torch.randn(1, 4, 3, 8, 8)
    for steps, shift in ((10, 16.0), (3, 5.0), (1, 3.0)):
        s = lr.FlowDPMSolverOracle()
        s.set_timesteps(steps, shift=shift)
        x = (1 - s.sigmas[0]) * x0 + s.sigmas[0] * n
        for _ in range(steps):
            x = s.step(n - x0, x)
        assert float((x - x0).abs().max()) < 5e-6


@pytest.mark.parametrize("steps,shift", [(8, 16.0), (3, 5.0), (2, 1.0), (1, 3.0)])
def test_dpm_host_coefficients_match_oracle(steps, shift):
    """FlowDPMSolverMultistepScheduler = the UniPC kernel with the corrector off: its per-step scalars pushed through the
    kernel's formula follow the independently written DPM-Solver++ oracle."""
    from versecrafter_b200 import schedulers
    sch = schedulers.FlowDPMSolverMultistepScheduler(shift=1)
    sch.set_timesteps(steps, shift=shift)
    orc = lr.FlowDPMSolverOracle()
    orc.set_timesteps(steps, shift=shift)
    assert torch.equal(sch.timesteps, orc.timesteps) and torch.equal(sch.sigmas, orc.sigmas)
    g = torch.Generator().manual_seed(steps + 100)
    x = torch.randn(1, 16, 2, 6, 8, generator=g).bfloat16()
    last = m0 = m1 = None
    sch._state = dict(have_last=False)
    for i in range(steps):
        u, c = (torch.randn(x.shape, generator=g).bfloat16() for _ in range(2))
        coef, op = sch._coefficients()
        assert int(coef[1]) == 0                                   # never a corrector
        out, x0, xc = _kernel_formula(coef, u, c, x, last, m0, m1, 5.0)
        m1, m0, last = m0, x0, xc
        sch._state["have_last"] = True
        sch._this_order = op
        sch._lower_order_nums = min(sch._lower_order_nums + 1, sch.solver_order)
        sch._step_index += 1
        want = orc.step(lr.cfg_combine(u, c, 5.0), x)
        x = out.bfloat16()
        d = (x.float() - want.float()).abs()
        assert float(d.max()) <= 2 ** -7 * max(1.0, float(want.float().abs().max())) and float((d > 0).float().mean()) < 0.01


@pytest.mark.gpu
def test_dpm_kernel_matches_oracle():
    from versecrafter_b200 import schedulers
    steps, shift, shape = 6, 12.0, (1, 16, 5, 30, 52)
    sch = schedulers.FlowDPMSolverMultistepScheduler(shift=1)
    sch.set_timesteps(steps, device="cuda", shift=shift)
    orc = lr.FlowDPMSolverOracle()
    orc.set_timesteps(steps, shift=shift)
    g = torch.Generator(device="cuda").manual_seed(41)
    x = torch.randn(shape, generator=g, device="cuda").bfloat16()
    for i in range(steps):
        u, c = (torch.randn(shape, generator=g, device="cuda").bfloat16() for _ in range(2))
        got = sch.step_cfg(u, c, x, 4.0)
        want = orc.step(lr.cfg_combine(u, c, 4.0).cpu(), x.cpu())
        d = (got.cpu().float() - want.float()).abs()
        assert float(d.max()) <= 2 ** -7 * max(1.0, float(want.float().abs().max())) and float((d > 0).float().mean()) < 0.01
        x = got
