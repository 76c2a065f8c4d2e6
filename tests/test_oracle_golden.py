"""Pins oracle/restate.py (fp32 mode) to fixtures produced by the reference's own code
(tests/golden/gen_golden.py).  CPU only."""
import ast
import os

import numpy as np
import pytest
import torch

from oracle import restate

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def relerr(a, b):
    a, b = torch.as_tensor(a).double(), torch.as_tensor(b).double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


@pytest.fixture(scope="module")
def pieces():
    return np.load(os.path.join(G, "pieces.npz"))


def test_rope_matches_reference(pieces):
    q = torch.from_numpy(pieces["rope_in"])
    grid = tuple(int(v) for v in pieces["rope_grid"])
    out = restate.rope_rotate(q, restate.rope_angles(q.shape[-1], grid))
    assert torch.equal(out, torch.from_numpy(pieces["rope_out"]))  # bit-exact, incl. pass-through pad rows


def test_sinusoid_matches_reference(pieces):
    out = restate.sinusoid_embedding(256, torch.from_numpy(pieces["sin_t"]))
    assert np.abs(out.numpy() - pieces["sin_out"]).max() < 1e-12


def test_norms_match_reference(pieces):
    x = torch.from_numpy(pieces["norm_in"])
    assert relerr(restate.rms_norm(x, torch.from_numpy(pieces["rms_gain"])), pieces["rms_out"]) < 1e-6
    ln = restate.layer_norm(x, torch.from_numpy(pieces["ln_w"]), torch.from_numpy(pieces["ln_b"]))
    assert relerr(ln, pieces["ln_out"]) < 1e-6


def test_patchify_and_unpatchify_match_reference(pieces):
    w, b, u = (torch.from_numpy(pieces[k]) for k in ("conv_w", "conv_b", "conv_in"))
    tok = restate.patch_embed(u, w, b, seq_len=int(pieces["conv_out"].shape[0]))[0]
    assert relerr(tok, pieces["conv_out"]) < 1e-5
    y = restate.unpatchify(torch.from_numpy(pieces["unp_in"])[0], (2, 2, 3))
    assert torch.equal(y, torch.from_numpy(pieces["unp_out"]))


@pytest.mark.parametrize("name", ["tiny4", "tiny3_layers013"])
def test_forward_matches_reference(name):
    g = np.load(os.path.join(G, f"forward_{name}.npz"))
    meta = ast.literal_eval(str(g["meta"]))
    cfg, grid = meta["cfg"], meta["grid"]
    sd = {k: v.to(torch.bfloat16).float() for k, v in
          restate.synth_state_dict(cfg, seed=meta["seed"], dtype=torch.float32).items()}
    inp = restate.synth_inputs(cfg, grid, seed=meta["seed"], t_value=meta["t"])
    from tests.golden.gen_golden import checksum
    assert abs(checksum(sd.values()) - float(g["weights_checksum"])) < 1e-6 * float(g["weights_checksum"]), \
        "torch CPU RNG stream differs from the one the fixtures were generated with"
    assert abs(checksum([inp["x"], inp["geoada_context"], *inp["context"]]) - float(g["inputs_checksum"])) \
        < 1e-6 * float(g["inputs_checksum"])
    rec = {}
    full_cfg = dict(cfg)
    y = restate.forward(sd, full_cfg, inp["x"], inp["t"], inp["geoada_context"], inp["context"], inp["seq_len"],
                        geoada_context_scale=meta["scale"], dtype=torch.float32, record=rec)
    for k in g.files:
        if k.startswith(("blk_out", "geo_out", "hint")):
            assert relerr(rec[k], g[k]) < 2e-5, k
    assert relerr(y, g["y"]) < 2e-5
    cos = torch.nn.functional.cosine_similarity(y.flatten().double(), torch.from_numpy(g["y"]).flatten().double(), dim=0)
    assert cos > 0.999999


def test_forward_matches_live_reference_when_present():
    """Extra pin with H=2 heads and a prompt-length mix, only where /root/reference exists."""
    from oracle import ref_import
    if not ref_import.reference_available():
        pytest.skip("reference tree not present on this machine")
    cfg = dict(dim=256, ffn_dim=256, num_heads=2, num_layers=2, text_dim=48, geoada_in_dim=128)
    _, vc = ref_import.load_reference_modules()
    model = vc.VerseCrafterWanTransformer3DModel(**cfg)
    sd = {k: v.to(torch.bfloat16).float() for k, v in restate.synth_state_dict(cfg, seed=9, dtype=torch.float32).items()}
    model.load_state_dict(sd, strict=True)
    model.eval()
    inp = restate.synth_inputs(cfg, (2, 4, 8), seed=9, t_value=999.0)
    with torch.no_grad():
        y_ref = model(x=inp["x"], t=inp["t"], geoada_context=inp["geoada_context"], context=inp["context"],
                      seq_len=inp["seq_len"], geoada_context_scale=0.5)
    y = restate.forward(sd, cfg, inp["x"], inp["t"], inp["geoada_context"], inp["context"], inp["seq_len"],
                        geoada_context_scale=0.5)
    assert relerr(y, y_ref) < 2e-5


def test_reference_key_mask_for_padded_sequences_cpu():
    """The attention() stand-in honours k_lens (wan_transformer3d.py:394-399): with seq_len > L the reference's own model gives
    the same latents as without padding (fp32, CPU), which un-masked attention over the zero-padded rows would not."""
    from oracle import ref_import
    from oracle.ref_run import run_reference
    if not ref_import.reference_available():
        pytest.skip("reference sources absent")
    mods = ref_import.load_reference_modules()
    cfg = dict(dim=128, ffn_dim=256, num_heads=1, num_layers=3, text_dim=32, geoada_in_dim=128, geoada_layers=[0, 1])
    a = run_reference(mods, cfg, (2, 6, 10), 5, 0.6, device="cpu", dtype=torch.float32)
    b = run_reference(mods, cfg, (2, 6, 10), 5, 0.6, device="cpu", dtype=torch.float32, seq_pad=9)
    assert b["cap"]["blk_in.0"].shape[1] == a["cap"]["blk_in.0"].shape[1] + 9
    assert torch.allclose(a["y"], b["y"], atol=2e-5, rtol=1e-4)
    from videox_fun.models import attention_utils
    q = torch.randn(2, 5, 1, 8)
    k, v = torch.randn(2, 7, 1, 8), torch.randn(2, 7, 1, 8)
    full = attention_utils.attention(q, k, v, k_lens=torch.tensor([7, 7]))
    cut = attention_utils.attention(q, k, v, k_lens=torch.tensor([7, 4]))
    assert torch.equal(full[0], cut[0]) and not torch.allclose(full[1], cut[1])
    assert torch.allclose(cut[1:], attention_utils.attention(q[1:], k[1:, :4], v[1:, :4]), atol=1e-6)
