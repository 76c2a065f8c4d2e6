"""Data formats on the input side of the control-map renderer: the loaders of the reference's
inference/rendering_4D_control_maps.py — `build_background` (:908-998), `load_camera_trajectory` (:1001-1009),
`load_ellipsoid_parameters` (:1012-1050) — with the same names, argument meaning and coordinate conventions, and
`render_scene()`, steps 1-9 of its `main()` (:1171-1330) on this library's kernels.  File decoding (PNG, mask resize and
dilation) goes through OpenCV exactly as in the reference; the arithmetic that builds the point cloud runs on the device
in plain torch (one pass over one 1280x720 frame, not a hot loop).  The demo scenes of the reference
(demo_data/*/camera_object_0) load through these functions unchanged."""
import json
from pathlib import Path

import numpy as np
import torch

from . import _lib

# OpenCV camera axes (x right, y down, z forward) -> Blender world axes (x right, y forward, z up)   (:59-63)
COORD_TRANSFORM_CV2BLENDER = np.array([[1, 0, 0], [0, 0, 1], [0, -1, 0]], dtype=np.float32)


def depth_to_3d(depth, intrinsic):
    """kornia.geometry.depth.depth_to_3d_v2(depth, K, normalize_points=False) (:969): pixel (u, v) with depth z ->
    z * K^-1 [u, v, 1]; depth [H, W], intrinsic [3, 3] in pixels -> [H, W, 3] fp32."""
    H, W = depth.shape
    dev = depth.device
    v, u = torch.meshgrid(torch.arange(H, device=dev, dtype=torch.float32), torch.arange(W, device=dev, dtype=torch.float32),
                          indexing="ij")
    fx, fy, cx, cy = intrinsic[0, 0], intrinsic[1, 1], intrinsic[0, 2], intrinsic[1, 2]
    x = (u - cx) / fx
    y = (v - cy) / fy
    return torch.stack([x, y, torch.ones_like(x)], dim=-1) * depth[..., None]


def build_background(png_path, npz_path, mask_dir, device="cuda"):
    """build_background (:908-998) -> (bg_points fp32 [P,3] in Blender world axes, bg_colors uint8 [P,3], intrinsic [3,3]
    in pixels, extrinsic [4,4] = identity, H, W).  The object masks (all PNGs of mask_dir, nearest-resized, OR-ed,
    dilated with a 10x10 ellipse) are cut out of the cloud; non-finite and |coordinate| >= 1e6 points are dropped."""
    import cv2
    image = cv2.imread(str(png_path))
    if image is None:
        raise _lib.VCError(f"build_background: cannot read {png_path}")
    image = cv2.cvtColor(image, cv2.COLOR_BGR2RGB)
    H, W = image.shape[:2]
    image_t = torch.from_numpy(image).to(device)
    data = np.load(npz_path)
    depth = torch.from_numpy(data["depth"].astype(np.float32)).to(device)
    intrinsic = torch.from_numpy(data["intrinsic"].astype(np.float32)).to(device).clone()
    intrinsic[0, 0] *= W   # the file stores normalised intrinsics (:936-941)
    intrinsic[1, 1] *= H
    intrinsic[0, 2] *= W
    intrinsic[1, 2] *= H
    extrinsic = torch.eye(4, dtype=torch.float32, device=device)
    combined = np.zeros((H, W), dtype=bool)
    for mask_file in sorted(Path(mask_dir).glob("*.png")) if mask_dir is not None else []:
        m = cv2.imread(str(mask_file), cv2.IMREAD_GRAYSCALE)
        m = cv2.resize(m, (W, H), interpolation=cv2.INTER_NEAREST)
        combined |= m > 127
    kernel = cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (10, 10))
    combined = cv2.dilate(combined.astype(np.uint8) * 255, kernel, iterations=1) > 127
    combined_t = torch.from_numpy(combined).to(device)
    pts_cam = depth_to_3d(depth, intrinsic).reshape(-1, 3)
    c2w = torch.linalg.inv(extrinsic)
    pts_world_cv = pts_cam @ c2w[:3, :3].T + c2w[:3, 3]
    pts_world = pts_world_cv @ torch.from_numpy(COORD_TRANSFORM_CV2BLENDER).to(device).T
    bg_mask = ~combined_t
    bg_points = pts_world[bg_mask.reshape(-1)]
    bg_colors = image_t[bg_mask]
    valid = torch.isfinite(bg_points).all(dim=1) & (bg_points.abs() < 1e6).all(dim=1)
    return bg_points[valid].contiguous(), bg_colors[valid].contiguous(), intrinsic, extrinsic, H, W


def load_camera_trajectory(trajectory_npz, device="cuda"):
    """load_camera_trajectory (:1001-1009): Blender camera-to-world matrices [N,4,4] -> OpenCV world-to-camera."""
    data = np.load(trajectory_npz)
    c2w = torch.from_numpy(data["extrinsics"].astype(np.float32)).to(device).clone()
    c2w[:, :3, 1:3] *= -1
    return torch.linalg.inv(c2w)


def load_ellipsoid_parameters(json_path, device="cuda"):
    """load_ellipsoid_parameters (:1012-1050) -> (gaussian_params_per_frame [{obj_id: (mean[3], cov[3,3])}],
    obj_id_to_color_idx, center_points_per_frame {frame: {obj_id: mean}})."""
    with open(json_path, "r") as f:
        data = json.load(f)
    obj_id_to_color_idx = dict(data["metadata"]["obj_id_to_color_idx"].items())
    per_frame, centers = [], {}
    for frame in data["frames"]:
        fi = frame["frame_index"]
        centers.setdefault(fi, {})
        params = {}
        for obj in frame["objects"]:
            oid = obj["object_id"]
            mean = torch.tensor(obj["gaussian_3d"]["mean"], dtype=torch.float32, device=device)
            cov = torch.tensor(obj["gaussian_3d"]["covariance"], dtype=torch.float32, device=device)
            params[oid] = (mean, cov)
            if obj["gaussian_3d"]["mean"] is not None:
                centers[fi][oid] = mean.clone()
        per_frame.append(params)
    return per_frame, obj_id_to_color_idx, centers


def render_scene(png_path, npz_path, mask_dir, trajectory_npz, ellipsoid_json, device="cuda", point_size=0.005,
                 ellipsoid_subdiv=3, gaussian_mask_threshold=0.003, ellipsoid_scale_factor=2.5, frames=None):
    """Steps 1-9 of the reference renderer's main() (:1171-1330) with its command-line defaults (:1146-1168): loads a
    scene in the reference's file formats and returns control_maps.render_control_maps()'s dict (the control videos as
    uint8 [N,H,W,3] CUDA tensors).  `frames`: optional list of frame indices (default: the whole trajectory; the depth
    range is then the range of those frames)."""
    from . import control_maps
    bg_points, bg_colors, K, _, H, W = build_background(png_path, npz_path, mask_dir, device)
    extrinsics = load_camera_trajectory(trajectory_npz, device)
    gp, color_idx, _ = load_ellipsoid_parameters(ellipsoid_json, device)
    n = len(extrinsics)
    sel = list(range(n)) if frames is None else list(frames)
    gp = [gp[f] if f < len(gp) else {} for f in sel]
    ell = [[(m, c) for m, c in g.values()] for g in gp]
    Ks = K.unsqueeze(0).repeat(len(sel), 1, 1)
    return control_maps.render_control_maps(bg_points, bg_colors, ell, gp, color_idx, Ks, extrinsics[sel], (H, W),
                                            point_size=point_size, gaussian_mask_threshold=gaussian_mask_threshold,
                                            scale_factor=ellipsoid_scale_factor, subdivisions=ellipsoid_subdiv)
