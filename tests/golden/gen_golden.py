"""Generates tests/golden/*.npz by running the REFERENCE's own code (imported from /root/reference through
oracle/ref_import.py).  Run in the build container only:  python tests/golden/gen_golden.py

Weights and inputs are NOT stored: they are regenerated from seeds by oracle.restate.synth_state_dict /
synth_inputs (pure torch CPU RNG); each fixture carries a float64 checksum of both so a drift of the RNG
stream is detected instead of silently comparing different models.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import ref_import, restate  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

CASES = {
    # name: (ctor cfg, latent (F,H,W), seed, t, scale)
    "tiny4": (dict(dim=256, ffn_dim=512, num_heads=2, num_layers=4, text_dim=64, geoada_in_dim=128),
              (3, 8, 12), 0, 500.0, 1.0),
    "tiny3_layers013": (dict(dim=128, ffn_dim=256, num_heads=1, num_layers=3, text_dim=32, geoada_in_dim=128,
                             geoada_layers=[0, 1]), (2, 6, 10), 5, 37.0, 0.6),
}


def checksum(tensors):
    return float(sum(float(t.double().abs().sum()) for t in tensors))


def run_case(name, cfg, grid, seed, t_value, scale):
    base, vc = ref_import.load_reference_modules()
    model = vc.VerseCrafterWanTransformer3DModel(**cfg)
    sd = restate.synth_state_dict(cfg, seed=seed, dtype=torch.float32)
    # weights are bf16-representable values held in fp32 (the reference then computes in fp32)
    sd = {k: v.to(torch.bfloat16).float() for k, v in sd.items()}
    missing, unexpected = model.load_state_dict(sd, strict=True), None
    model.eval().requires_grad_(False)
    inp = restate.synth_inputs(cfg, grid, seed=seed, t_value=t_value)

    rec = {}
    hooks = []
    for i, blk in enumerate(model.blocks):
        hooks.append(blk.register_forward_hook(lambda m, a, out, i=i: rec.__setitem__(f"blk_out.{i}", out.clone())))
    for j, blk in enumerate(model.geoada_blocks):
        # output is the stacked tensor [hint_0 .. hint_j, c]  (wan_transformer3d_versecrafter.py:121-124)
        def hook(m, a, out, j=j):
            rec[f"geo_out.{j}"] = out[-1].clone()
            rec[f"hint.{j}"] = out[-2].clone()
        hooks.append(blk.register_forward_hook(hook))
    with torch.no_grad():
        y = model(x=inp["x"], t=inp["t"], geoada_context=inp["geoada_context"], context=inp["context"],
                  seq_len=inp["seq_len"], geoada_context_scale=scale)
    for h in hooks:
        h.remove()
    out = {k: v.numpy().astype(np.float32) for k, v in rec.items()}
    out["y"] = y.numpy().astype(np.float32)
    out["weights_checksum"] = np.float64(checksum(sd.values()))
    out["inputs_checksum"] = np.float64(checksum([inp["x"], inp["geoada_context"], *inp["context"]]))
    out["meta"] = np.array(repr(dict(cfg=cfg, grid=grid, seed=seed, t=t_value, scale=scale)))
    np.savez_compressed(os.path.join(HERE, f"forward_{name}.npz"), **out)
    print(name, "y mean|.|", float(np.abs(out["y"]).mean()), "keys", len(out))


def run_pieces():
    base, _ = ref_import.load_reference_modules()
    g = torch.Generator().manual_seed(3)
    out = {}
    # RoPE: rope_apply (wan_transformer3d.py:143-172) with the model's freqs table (:783-795)
    d, H, grid = 128, 2, (3, 4, 6)
    freqs = torch.cat([base.rope_params(1024, d - 4 * (d // 6)), base.rope_params(1024, 2 * (d // 6)),
                       base.rope_params(1024, 2 * (d // 6))], dim=1)
    L = grid[0] * grid[1] * grid[2]
    q = torch.randn(2, L + 5, H, d, generator=g)  # 5 pad rows pass through
    out["rope_in"] = q.numpy()
    out["rope_out"] = base.rope_apply(q, torch.tensor([grid, grid]), freqs).numpy()
    out["rope_grid"] = np.array(grid)
    # sinusoid (:39-49)
    t = torch.tensor([0.0, 20.0, 500.0, 999.0])
    out["sin_t"] = t.numpy()
    out["sin_out"] = base.sinusoidal_embedding_1d(256, t).numpy()
    # RMSNorm / LayerNorm modules (:307-346)
    x = torch.randn(3, 7, 256, generator=g)
    rn = base.WanRMSNorm(256, eps=1e-6)
    rn.weight.data = 1 + 0.1 * torch.randn(256, generator=g)
    out["norm_in"] = x.numpy()
    out["rms_gain"] = rn.weight.data.numpy()
    out["rms_out"] = rn(x).detach().numpy()
    ln = base.WanLayerNorm(256, eps=1e-6, elementwise_affine=True)
    ln.weight.data = 1 + 0.1 * torch.randn(256, generator=g)
    ln.bias.data = 0.02 * torch.randn(256, generator=g)
    out["ln_w"], out["ln_b"] = ln.weight.data.numpy(), ln.bias.data.numpy()
    out["ln_out"] = ln(x).detach().numpy()
    # Conv3d patch embedding vs patchify-GEMM, and unpatchify (:1127-1150)
    conv = torch.nn.Conv3d(6, 32, kernel_size=(1, 2, 2), stride=(1, 2, 2))
    u = torch.randn(6, 2, 4, 6, generator=g)
    out["conv_w"], out["conv_b"], out["conv_in"] = conv.weight.data.numpy(), conv.bias.data.numpy(), u.numpy()
    out["conv_out"] = conv(u[None]).flatten(2).transpose(1, 2)[0].detach().numpy()

    class _U:  # unpatchify only touches out_dim / patch_size
        out_dim, patch_size = 16, (1, 2, 2)
    yy = torch.randn(1, 2 * 2 * 3, 64, generator=g)
    out["unp_in"] = yy.numpy()
    out["unp_out"] = base.WanTransformer3DModel.unpatchify(_U(), yy, torch.tensor([[2, 2, 3]]))[0].numpy()
    np.savez_compressed(os.path.join(HERE, "pieces.npz"), **out)
    print("pieces", list(out))


if __name__ == "__main__":
    torch.set_num_threads(8)
    run_pieces()
    for name, (cfg, grid, seed, t_value, scale) in CASES.items():
        run_case(name, cfg, grid, seed, t_value, scale)
