"""Generates tests/golden/control_maps.npz by executing the REFERENCE's own per-pixel control-map functions
(/root/reference/inference/rendering_4D_control_maps.py).  The module cannot be imported here (pytorch3d, kornia,
matplotlib are absent), so the function definitions are cut out of the unmodified source with `ast` and executed with
torch / numpy in scope and a stand-in `matplotlib.colormaps['tab20']` (oracle.control_maps_restate.tab20).
Run in the build container only:  python tests/golden/gen_golden_control_maps.py"""
import ast
import logging
import os
import sys
import textwrap
from types import SimpleNamespace
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import control_maps_restate as cr  # noqa: E402

SRC = "/root/reference/inference/rendering_4D_control_maps.py"
HERE = os.path.dirname(os.path.abspath(__file__))
NAMES = ["composite_by_depth", "merge_bg_and_fg_mask", "visualize_depth_as_grayscale", "project_3d_gaussians_to_2d",
         "blend_gaussian_projection_with_bg", "project_gaussian_to_2d_gpu", "compute_probability_density_map_gpu",
         "get_object_color"]


def reference_functions():
    src = open(SRC).read()
    fake_mpl = SimpleNamespace(colormaps={"tab20": lambda i: cr.tab20(i) + (1.0,)})
    ns = {"torch": torch, "np": np, "logger": logging.getLogger("ref"), "matplotlib": fake_mpl, "Dict": Dict, "List": List,
          "Optional": Optional, "Tuple": Tuple}
    for node in ast.parse(src).body:
        if isinstance(node, ast.FunctionDef) and node.name in NAMES:
            exec(compile(textwrap.dedent(ast.get_source_segment(src, node)), SRC, "exec"), ns)
    return SimpleNamespace(**{n: ns[n] for n in NAMES})


def scene(seed, H, W, n_frames):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.rand(*s, generator=g)
    bg_rgb = (r(n_frames, H, W, 3) * 255).to(torch.uint8)
    fg_rgb = (r(n_frames, H, W, 3) * 255).to(torch.uint8)
    bg_depth = torch.where(r(n_frames, H, W) > 0.2, 0.5 + 4 * r(n_frames, H, W), torch.zeros(n_frames, H, W))
    fg_depth = torch.where(r(n_frames, H, W) > 0.5, 0.5 + 4 * r(n_frames, H, W), torch.zeros(n_frames, H, W))
    fg_depth[0, :2] = bg_depth[0, :2]                    # ties and near-ties at the 1e-6 margin
    fg_depth[0, 2:4] = bg_depth[0, 2:4] - 2e-6
    bg_mask = bg_depth > 0
    fg_mask = (fg_depth > 0) | (r(n_frames, H, W) > 0.9)
    return bg_rgb, bg_depth, bg_mask, fg_rgb, fg_depth, fg_mask


def gaussians(seed, n_frames, n_obj):
    g = torch.Generator().manual_seed(seed)
    frames = []
    for f in range(n_frames):
        d = {}
        for o in range(n_obj):
            mean = torch.tensor([-0.8 + 0.8 * o + 0.05 * f, 0.1 * o - 0.1, 2.0 + 0.7 * o - 0.1 * f])
            A = torch.randn(3, 3, generator=g) * 0.15
            d[str(o + 1)] = (mean, A @ A.T + 0.01 * torch.eye(3))
        if f == 1:
            d["9"] = (torch.tensor([0.0, 0.0, 0.3]), 0.01 * torch.eye(3))    # z <= 0.5: skipped density, still listed
            d["8"] = (torch.tensor([0.0, 0.0, -1.0]), 0.01 * torch.eye(3))   # behind the camera: dropped
        frames.append(d)
    return frames


def main():
    ref = reference_functions()
    out = {}
    H, W, N = 36, 64, 3
    bg_rgb, bg_depth, bg_mask, fg_rgb, fg_depth, fg_mask = scene(1, H, W, N)
    comp = [ref.composite_by_depth(bg_rgb[i], bg_depth[i], fg_rgb[i], fg_depth[i], fg_mask[i]) for i in range(N)]
    out["comp_rgb"] = torch.stack([c[0] for c in comp]).numpy()
    out["comp_depth"] = torch.stack([c[1] for c in comp]).numpy()
    out["merged_mask"] = torch.stack(ref.merge_bg_and_fg_mask(list(bg_depth), list(fg_depth), list(bg_mask), list(fg_mask),
                                                              device="cpu")).numpy()
    for tag, (lo, hi) in {"norm": (0.62, 4.3), "raw": (0.0, 1.0)}.items():
        out[f"gray_{tag}"] = torch.stack(ref.visualize_depth_as_grayscale(list(bg_depth), lo, hi)).numpy()
        out[f"gray_{tag}_range"] = np.array([lo, hi])
    K = np.array([[60.0, 0, W / 2], [0, 60.0, H / 2], [0, 0, 1]], dtype=np.float32)
    Ts = []
    for f in range(N):
        T = np.eye(4, dtype=np.float32)
        T[:3, 3] = [0.05 * f, -0.02 * f, 0.1 * f]
        Ts.append(T)
    gp = gaussians(7, N, 3)
    idx = {"1": 0, "2": 3, "3": 6, "8": 1, "9": 2}
    rgb, alpha = ref.project_3d_gaussians_to_2d(gp, idx, [K] * N, Ts, (W, H), threshold=0.003, device="cpu")
    out["g_rgb"], out["g_alpha"] = torch.stack(rgb).numpy(), torch.stack(alpha).numpy()
    out["g_blend"] = torch.stack(ref.blend_gaussian_projection_with_bg(rgb, alpha, list(bg_rgb))).numpy()
    out["scene_seed"], out["gauss_seed"], out["shape"] = np.array(1), np.array(7), np.array([H, W, N])
    out["K"], out["T"] = K, np.stack(Ts)
    np.savez_compressed(os.path.join(HERE, "control_maps.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
