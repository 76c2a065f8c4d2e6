"""Generates tests/golden/loop_pieces.npz by executing the REFERENCE's own `geoada_encode_masks` and `geoada_latent`
(methods of WanVerseCrafterPipeline; the file cannot be imported here because diffusers / videox_fun are absent, so the
two function definitions are cut out of the unmodified source with `ast` and executed with torch / F in scope).
Run in the build container only:  python tests/golden/gen_golden_loop.py"""
import ast
import os

import numpy as np
import torch
import torch.nn.functional as F

SRC = "/root/reference/versecrafter/pipeline/pipeline_wan_versecrafter.py"
HERE = os.path.dirname(os.path.abspath(__file__))


def reference_functions(names):
    src = open(SRC).read()
    tree = ast.parse(src)
    ns = {"torch": torch, "F": F}
    for node in ast.walk(tree):
        if isinstance(node, ast.FunctionDef) and node.name in names:
            code = ast.get_source_segment(src, node)
            exec(compile(ast.Module(body=[ast.parse(__import__("textwrap").dedent(code)).body[0]], type_ignores=[]),
                         SRC, "exec"), ns)
    return [ns[n] for n in names]


def main():
    enc, cat = reference_functions(["geoada_encode_masks", "geoada_latent"])
    g = torch.Generator().manual_seed(123)
    out = {}
    for name, (T, H, W) in {"a": (9, 32, 48), "b": (81, 16, 32), "c": (13, 48, 16)}.items():
        mask = (torch.rand(3, T, H, W, generator=g) > 0.5).float() * torch.rand(3, T, H, W, generator=g)
        m = enc(None, [mask])[0]                       # self is unused by the method
        lat = [torch.randn(16, m.shape[1], m.shape[2], m.shape[3], generator=g) for _ in range(4)]
        ctx = cat(None, [torch.cat(lat, dim=0)], [m])[0]
        out[f"{name}_mask"] = mask.numpy()
        out[f"{name}_lat"] = torch.stack(lat).numpy()
        out[f"{name}_enc"] = m.numpy()
        out[f"{name}_ctx"] = ctx.numpy()
        refs = [torch.zeros(3, 1, H, W), torch.zeros(3, 1, H, W)]       # two reference images -> 2 zero frames
        out[f"{name}_enc_ref2"] = enc(None, [mask], [refs])[0].numpy()
    np.savez_compressed(os.path.join(HERE, "loop_pieces.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
