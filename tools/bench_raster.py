"""Timing of the N4 control-map kernels at the reference's working size (81 frames, 1280x720, ~0.9 M background points,
3 ellipsoids of 1280 faces) on one B200: CUDA events around whole frame stacks."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from versecrafter_b200 import control_maps as cm, raster  # noqa: E402


def timed(fn, iters=2):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters, out


def main():
    H, W, N = 720, 1280, 81
    g = torch.Generator(device="cuda").manual_seed(0)
    K = np.array([[900.0, 0, W / 2], [0, 900.0, H / 2], [0, 0, 1]], dtype=np.float32)
    # background cloud: one point per pixel of a depth map (build_background :908-1000), camera dollying forward
    v, u = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
    z = 2.0 + 3.0 * torch.rand(H, W, generator=g, device="cuda")
    pts = torch.stack([(u + 0.5 - W / 2) / 900.0 * z, (v + 0.5 - H / 2) / 900.0 * z, z], -1).reshape(-1, 3).float()
    col = (torch.rand(H * W, 3, generator=g, device="cuda") * 255).to(torch.uint8)
    Ts = []
    for f in range(N):
        T = np.eye(4, dtype=np.float32)
        T[:3, 3] = [0.002 * f, 0.0, -0.01 * f]
        Ts.append(T)
    ms, (rgb, depth, mask) = timed(lambda: raster.render_point_cloud_batch(pts, col, [K] * N, Ts, (H, W), point_size=0.005), 1)
    out = {"points": int(pts.shape[0]), "frames": N, "size": [H, W], "render_points_ms": ms,
           "render_points_ms_per_frame": ms / N, "coverage": float(mask.float().mean())}
    meshes = []
    for f in range(N):
        parts = [raster.make_ellipsoid_mesh([-0.6 + 0.6 * o, 0.05 * o, 3.0 + 0.3 * o - 0.005 * f],
                                            np.diag([0.05, 0.03, 0.04]).astype(np.float32), 2.5, 3) for o in range(3)]
        meshes.append(raster.combine_meshes_for_scene(parts))
    ms, (fdepth, fmask) = timed(lambda: raster.render_meshes_depth_batch(meshes, [K] * N, Ts, (H, W)))
    out.update(mesh_depth_ms=ms, mesh_coverage=float(fmask.float().mean()))
    fg_rgb = torch.zeros_like(rgb)
    ms, _ = timed(lambda: cm.composite_by_depth(rgb, depth, fg_rgb, fdepth, fmask, bg_mask=mask))
    out["composite_ms"] = ms
    out["composite_gbs"] = N * H * W * (3 + 4 + 1 + 3 + 4 + 1 + 3 + 4 + 3) / ms / 1e6
    ms, _ = timed(lambda: cm.visualize_depth_as_grayscale(depth, 2.0, 5.0))
    out["depth_gray_ms"] = ms
    out["depth_gray_gbs"] = N * H * W * 7 / ms / 1e6
    gp = [{str(o): (torch.tensor([-0.6 + 0.6 * o, 0.05 * o, 3.0 + 0.3 * o - 0.005 * f]), torch.diag(torch.tensor([0.05, 0.03, 0.04])))
           for o in range(3)} for f in range(N)]
    ms, _ = timed(lambda: cm.project_3d_gaussians_to_2d(gp, {"0": 0, "1": 2, "2": 4}, [K] * N, Ts, (W, H), threshold=0.003,
                                                        background_frames=rgb), 1)
    out["gaussian_projection_ms"] = ms
    print(json.dumps(out), flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/bench_raster.json", "w") as fh:
        fh.write(json.dumps(out) + "\n")


if __name__ == "__main__":
    main()
