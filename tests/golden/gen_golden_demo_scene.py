"""Assembles tests/golden/demo_scene/: one of the reference's own demo scenes (category (b) fixtures, copied byte for byte
from /root/reference/demo_data/<scene>/) in the reference's file formats —
    0001.png, depth_intrinsics.npz (MoGe depth + normalised intrinsics), masks/*.png (Grounded-SAM-2 object masks),
    custom_camera_trajectory.npz, custom_3D_gaussian_trajectory.json (Blender export)
— and, under reference_maps/, the four control videos the REFERENCE ITSELF rendered for that scene with its pytorch3d
renderer (inference/rendering_4D_control_maps.py with its command-line defaults, inference.sh step 5): background_depth,
3D_gaussian_depth, merged_mask, 3D_gaussian_RGB (h264, 81 frames of 1280x720).  They are the golden outputs that pin the
rasteriser restatement (oracle/raster_restate.py, oracle/raster_fast.py) and the control-map chain: tests/test_demo_scene.py.
Also writes oracle_depth_range.json: the global depth range (0.1 % / 99 % quantiles over the three depth stacks of all 81
frames, reference :541-571) computed by the oracle chain, so that the CPU test needs to render a few frames only.

    python tests/golden/gen_golden_demo_scene.py          (needs /root/reference; ~3 min of CPU)
"""
import json
import os
import shutil
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
SCENE = "dc58debb-191d-50cb-8ba1-a5afdeec1808_0000091_0000172"
SRC = os.path.join(os.environ.get("VC_REFERENCE_ROOT", "/root/reference"), "demo_data", SCENE)
DST = os.path.join(HERE, "demo_scene")
FILES = {
    "0001.png": "0001.png",
    "estimated_depth/depth_intrinsics.npz": "depth_intrinsics.npz",
    "camera_object_0/custom_camera_trajectory.npz": "custom_camera_trajectory.npz",
    "camera_object_0/custom_3D_gaussian_trajectory.json": "custom_3D_gaussian_trajectory.json",
    "camera_object_0/rendering_4D_maps/background_depth.mp4": "reference_maps/background_depth.mp4",
    "camera_object_0/rendering_4D_maps/3D_gaussian_depth.mp4": "reference_maps/3D_gaussian_depth.mp4",
    "camera_object_0/rendering_4D_maps/merged_mask.mp4": "reference_maps/merged_mask.mp4",
    "camera_object_0/rendering_4D_maps/3D_gaussian_RGB.mp4": "reference_maps/3D_gaussian_RGB.mp4",
}


def oracle_depth_stacks(frames=None):
    """(bg_depth, fg_depth, bg_mask, fg_mask) [N,H,W] torch tensors of the oracle chain on the committed scene."""
    from oracle import raster_fast as rf, raster_restate as rr
    from versecrafter_b200 import scene_io
    pts, _, K, _, H, W = scene_io.build_background(f"{DST}/0001.png", f"{DST}/depth_intrinsics.npz", f"{DST}/masks", "cpu")
    ext = scene_io.load_camera_trajectory(f"{DST}/custom_camera_trajectory.npz", "cpu")
    gp, _, _ = scene_io.load_ellipsoid_parameters(f"{DST}/custom_3D_gaussian_trajectory.json", "cpu")
    out = [[], [], [], []]
    for f in (range(len(ext)) if frames is None else frames):
        bd, bm = rf.render_points_depth(pts.numpy(), K.numpy(), ext[f].numpy(), (H, W), 0.005)
        meshes = [rr.ellipsoid_mesh(m.numpy(), c.numpy(), scale_factor=2.5, subdivisions=3) for m, c in gp[f].values()]
        verts = np.concatenate([v for v, _ in meshes])
        faces = np.concatenate([fc + sum(len(v) for v, _ in meshes[:i]) for i, (_, fc) in enumerate(meshes)])
        fd, fm = rf.render_mesh_depth(verts, faces, K.numpy(), ext[f].numpy(), (H, W))
        for lst, a in zip(out, (bd, fd, bm, fm)):
            lst.append(torch.from_numpy(a))
    return [torch.stack(x) for x in out]


if __name__ == "__main__":
    for rel, dst in FILES.items():
        os.makedirs(os.path.dirname(os.path.join(DST, dst)), exist_ok=True)
        shutil.copyfile(os.path.join(SRC, rel), os.path.join(DST, dst))
    os.makedirs(os.path.join(DST, "masks"), exist_ok=True)
    for n in sorted(os.listdir(os.path.join(SRC, "object_mask", "masks"))):
        shutil.copyfile(os.path.join(SRC, "object_mask", "masks", n), os.path.join(DST, "masks", n))
    from oracle import control_maps_restate as cr
    from versecrafter_b200 import control_maps as cm
    BD, FD, BM, FM = oracle_depth_stacks()
    z = torch.zeros(*BD.shape, 3, dtype=torch.uint8)
    _, comb = cr.composite_by_depth(z, BD, z, FD, FM)
    lo, hi = cm.compute_global_depth_range([BD, FD, comb], generator=torch.Generator().manual_seed(0))
    with open(os.path.join(DST, "oracle_depth_range.json"), "w") as f:
        json.dump({"min_depth": lo, "max_depth": hi, "frames": int(BD.shape[0]), "scene": SCENE,
                   "note": "quantiles of a random 1M-value subset (torch.Generator seed 0), as the reference subsamples"}, f, indent=1)
    print("depth range", lo, hi)
