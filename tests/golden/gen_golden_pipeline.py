"""Generates tests/golden/pipeline_tiny.npz: the latents after 3 steps of the reference's own sampling pipeline
(WanVerseCrafterPipeline.__call__, imported unmodified through oracle/ref_pipeline.py) around the reference's own
transformer in fp32 on CPU, tiny config, with the fake VAE / text encoder / Euler scheduler of oracle/ref_pipeline.py.

    python tests/golden/gen_golden_pipeline.py        (needs /root/reference or oracle/_ref)
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import ref_import, ref_pipeline, restate  # noqa: E402

CFG = dict(dim=256, ffn_dim=512, num_heads=2, num_layers=2, text_dim=64, geoada_in_dim=128)
SEED, STEPS, SCALE, GUIDANCE = 21, 3, 0.8, 6.0


def reference_model(device="cpu", dtype=torch.float32):
    _, vc = ref_import.load_reference_modules()
    sd = restate.synth_state_dict(CFG, seed=SEED, device=device, dtype=dtype)
    ref = vc.VerseCrafterWanTransformer3DModel(**CFG)
    ref.load_state_dict(sd, strict=True)
    return ref.to(device=device, dtype=dtype).eval().requires_grad_(False), sd


if __name__ == "__main__":
    ref, _ = reference_model()
    calls = []
    lat = ref_pipeline.run_pipeline(ref, CFG["text_dim"], "cpu", torch.float32, steps=STEPS, guidance_scale=GUIDANCE,
                                    geoada_context_scale=SCALE, calls=calls)
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "pipeline_tiny.npz")
    np.savez_compressed(out, latents=lat.numpy(), meta=str(dict(cfg=CFG, seed=SEED, steps=STEPS, scale=SCALE, guidance=GUIDANCE,
                                                                 calls=len(calls))))
    print(out, lat.shape, float(lat.abs().mean()))
