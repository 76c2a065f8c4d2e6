"""Per-pixel control-map stages of the reference's 4D control-map renderer as kernels of libvc_b200.so
(/root/reference/inference/rendering_4D_control_maps.py).  Function names and argument meaning
follow the reference; tensors are CUDA tensors (uint8 images [.., H, W, 3], fp32 depths, bool masks), whole frame stacks
are processed in one launch.  The pytorch3d point-cloud / mesh rasterisers that produce the inputs are not part of this
module.  No CPU fallback: CPU tensors raise VCError."""
import ctypes

import torch

from . import _lib, ops

TAB20_255 = [(31, 119, 180), (174, 199, 232), (255, 127, 14), (255, 187, 120), (44, 160, 44), (152, 223, 138),
             (214, 39, 40), (255, 152, 150), (148, 103, 189), (197, 176, 213), (140, 86, 75), (196, 156, 148),
             (227, 119, 194), (247, 182, 210), (127, 127, 127), (199, 199, 199), (188, 189, 34), (219, 219, 141),
             (23, 190, 207), (158, 218, 229)]   # matplotlib 'tab20' (get_object_color :885-907)


def _cu(t, dtype, name):
    if not t.is_cuda:
        raise _lib.VCError(f"{name}: expected a CUDA tensor (no CPU fallback)")
    if t.dtype == torch.bool and dtype == torch.uint8:
        t = t.view(torch.uint8)
    if t.dtype != dtype:
        raise _lib.VCError(f"{name}: expected {dtype}, got {t.dtype}")
    return t.contiguous()


def get_object_color(obj_id, obj_id_to_color_idx, return_float=False):
    c = TAB20_255[obj_id_to_color_idx.get(obj_id, 0) % 20]
    return tuple(v / 255.0 for v in c) if return_float else c


def composite_by_depth(bg_rgb, bg_depth, fg_rgb, fg_depth, fg_mask, bg_mask=None):
    """composite_by_depth / composite_by_depth_batch (:398-411, :438-455) -> (out_rgb, out_depth); with `bg_mask` also
    the merged mask frames of merge_bg_and_fg_mask (:736-763) as a third result, from the same pass."""
    bg_rgb, fg_rgb = _cu(bg_rgb, torch.uint8, "bg_rgb"), _cu(fg_rgb, torch.uint8, "fg_rgb")
    bg_depth, fg_depth = _cu(bg_depth, torch.float32, "bg_depth"), _cu(fg_depth, torch.float32, "fg_depth")
    fg_mask = _cu(fg_mask, torch.uint8, "fg_mask")
    if not (bg_depth.shape == fg_depth.shape == fg_mask.shape and bg_rgb.shape == fg_rgb.shape == bg_depth.shape + (3,)):
        raise _lib.VCError("composite_by_depth: shape mismatch")
    out_rgb, out_depth = torch.empty_like(bg_rgb), torch.empty_like(bg_depth)
    out_mask = None
    if bg_mask is not None:
        bg_mask = _cu(bg_mask, torch.uint8, "bg_mask")
        out_mask = torch.empty_like(bg_rgb)
    lib = _lib.load()
    _lib.check(lib.vc_composite_by_depth(ops._ptr(bg_rgb), ops._ptr(bg_depth), ops._ptr(bg_mask), ops._ptr(fg_rgb),
                                         ops._ptr(fg_depth), ops._ptr(fg_mask), ops._ptr(out_rgb), ops._ptr(out_depth),
                                         ops._ptr(out_mask), bg_depth.numel(), ops._stream()), "vc_composite_by_depth")
    ops.launch_count += 1
    return (out_rgb, out_depth) if out_mask is None else (out_rgb, out_depth, out_mask)


def merge_bg_and_fg_mask(bg_depth, fg_depth, bg_mask, fg_mask):
    """merge_bg_and_fg_mask (:736-763) on a frame stack -> uint8 [.., H, W, 3]."""
    bg_depth, fg_depth = _cu(bg_depth, torch.float32, "bg_depth"), _cu(fg_depth, torch.float32, "fg_depth")
    bg_mask, fg_mask = _cu(bg_mask, torch.uint8, "bg_mask"), _cu(fg_mask, torch.uint8, "fg_mask")
    out = torch.empty(bg_depth.shape + (3,), dtype=torch.uint8, device=bg_depth.device)
    lib = _lib.load()
    _lib.check(lib.vc_composite_by_depth(None, ops._ptr(bg_depth), ops._ptr(bg_mask), None, ops._ptr(fg_depth),
                                         ops._ptr(fg_mask), None, None, ops._ptr(out), bg_depth.numel(), ops._stream()),
               "vc_composite_by_depth")
    ops.launch_count += 1
    return out


def visualize_depth_as_grayscale(depth, global_min_depth, global_max_depth):
    """visualize_depth_as_grayscale (:487-539) with the global range of compute_global_depth_range (python floats)."""
    depth = _cu(depth, torch.float32, "depth")
    lo, hi = float(global_min_depth), float(global_max_depth)
    normalise = hi > 0 and lo > 0
    min_disp = 1.0 / hi if normalise else 0.0
    denom = (1.0 / lo - min_disp + 1e-8) if normalise else 1.0
    out = torch.empty(depth.shape + (3,), dtype=torch.uint8, device=depth.device)
    lib = _lib.load()
    _lib.check(lib.vc_depth_to_gray(ops._ptr(depth), ops._ptr(out), depth.numel(), min_disp, denom, int(normalise),
                                    ops._stream()), "vc_depth_to_gray")
    ops.launch_count += 1
    return out


def _project_gaussian(mean, cov, K, R, t):
    """2-D mean / inverse covariance / pdf coefficient of one Gaussian, fp32 on the host in the reference's op order
    (compute_probability_density_map_gpu :781-864).  None when the reference skips the Gaussian."""
    m_cam = R @ mean + t
    c_cam = R @ cov @ R.T
    x, y, z = m_cam[0], m_cam[1], m_cam[2]
    if z <= 0.5:
        return None
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    J = torch.tensor([[fx / z, 0, -fx * x / (z * z)], [0, fy / z, -fy * y / (z * z)]], dtype=torch.float32)
    mean_2d = torch.tensor([fx * x / z + cx, fy * y / z + cy], dtype=torch.float32)
    cov_2d = J @ c_cam @ J.T
    cov_2d += torch.eye(2) * 1e-6
    det = torch.det(cov_2d)
    if det > 1e11:
        return None
    inv = torch.linalg.inv(cov_2d)
    coeff = 1.0 / (2 * torch.pi * torch.sqrt(det))
    return [float(mean_2d[0]), float(mean_2d[1]), float(inv[0, 0]), float(inv[0, 1]), float(inv[1, 0]), float(inv[1, 1]),
            float(coeff)]


def project_3d_gaussians_to_2d(gaussian_params_per_frame, obj_id_to_color_idx, intrinsics, extrinsics, image_size,
                               threshold=0.05, device="cuda", background_frames=None):
    """project_3d_gaussians_to_2d (:573-694) -> (rgb_frames uint8 [N,H,W,3], alpha_frames fp32 [N,H,W]); with
    `background_frames` (uint8 [N,H,W,3]) also blend_gaussian_projection_with_bg (:697-733) and main()'s masked
    projection frames: returns (rgb, alpha, blended, masked).  Per-Gaussian 3x3 projection algebra runs on the host
    (a handful of scalars per frame); every per-pixel operation runs in the kernels."""
    W, H = image_size
    n = min(len(gaussian_params_per_frame), len(intrinsics), len(extrinsics))
    rgb = torch.empty((n, H, W, 3), dtype=torch.uint8, device=device)
    alpha = torch.empty((n, H, W), dtype=torch.float32, device=device)
    blend = masked = None
    if background_frames is not None:
        background_frames = _cu(background_frames, torch.uint8, "background_frames")
        blend, masked = torch.empty_like(rgb), torch.empty_like(rgb)
    scratch = torch.zeros(32, dtype=torch.int32, device=device)
    lib = _lib.load()
    for f in range(n):
        K = torch.as_tensor(intrinsics[f]).float().cpu()
        T = torch.as_tensor(extrinsics[f]).float().cpu()
        R, t = T[:3, :3], T[:3, 3]
        items = []
        for obj_id, (mean, cov) in gaussian_params_per_frame[f].items():
            mean, cov = torch.as_tensor(mean).float().cpu(), torch.as_tensor(cov).float().cpu()
            depth_z = float((R @ mean + t)[2])
            if depth_z > 0:                                            # :650
                p = _project_gaussian(mean, cov, K, R, t) or [0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0]   # skipped: zero density
                items.append((depth_z, p + list(get_object_color(obj_id, obj_id_to_color_idx, return_float=True))))
        items.sort(key=lambda it: it[0], reverse=True)                 # far -> near (:668); stable like list.sort
        flat = [v for _, p in items for v in p]
        arr = (ctypes.c_float * max(len(flat), 1))(*flat)
        _lib.check(lib.vc_gaussian_projection(len(items), arr, W, H, float(threshold),
                                              ops._ptr(background_frames[f]) if blend is not None else None,
                                              ops._ptr(scratch), ops._ptr(rgb[f]), ops._ptr(alpha[f]),
                                              ops._ptr(blend[f]) if blend is not None else None,
                                              ops._ptr(masked[f]) if masked is not None else None, ops._stream()),
                   "vc_gaussian_projection")
        ops.launch_count += 2
    return (rgb, alpha) if blend is None else (rgb, alpha, blend, masked)


def compute_global_depth_range(depth_stacks, max_samples=1000000, generator=None):
    """compute_global_depth_range (:541-571): 0.1 % / 99 % quantiles of all positive depths of the given frame stacks
    (a random subset of `max_samples` values when there are more, as in the reference) -> (min_depth, max_depth) python
    floats; (0.0, 1.0) when nothing is valid.  Plain torch on the stacks' device: a reduction over a scalar range, not a
    per-pixel stage."""
    vals = [d[d > 0].flatten() for d in depth_stacks if d.numel() and bool((d > 0).any())]
    if not vals:
        return 0.0, 1.0
    v = torch.cat(vals)
    if v.numel() > max_samples:
        v = v[torch.randperm(v.numel(), device=v.device, generator=generator)[:max_samples]]
    return float(torch.quantile(v, 0.001)), float(torch.quantile(v, 0.99))


def render_control_maps(background_points, background_colors, ellipsoids_per_frame, gaussian_params_per_frame,
                        obj_id_to_color_idx, intrinsics, extrinsics, image_size, point_size=0.01,
                        gaussian_mask_threshold=0.05, depth_range=None, scale_factor=2.0, subdivisions=3):
    """Steps 5-9 of the reference renderer's main() (:1240-1330) on this library's kernels, whole frame stacks at once:
    background point-cloud splat, ellipsoid depth/mask, global-range depth -> gray, merged mask, 3D-Gaussian RGB
    projection, and step 10's preview composite (:1350-1373, "background_and_3D_gaussian").  Returns the five control videos
    the VerseCrafter pipeline consumes and the preview as uint8 [N, H, W, 3] CUDA tensors (names = the reference's output
    files) plus the intermediate depths/masks.
      background_points fp32 [P,3], background_colors uint8 [P,3] (CUDA); ellipsoids_per_frame: per frame a list of
      (mean[3], cov[3,3]) (may be empty); gaussian_params_per_frame: per frame {obj_id: (mean, cov)};
      intrinsics [N,3,3], extrinsics [N,4,4]; image_size = (H, W)."""
    from . import raster
    H, W = image_size
    N = len(extrinsics)
    dev = background_points.device
    bg_rgb, bg_depth, bg_mask = raster.render_point_cloud_batch(background_points, background_colors, intrinsics, extrinsics,
                                                                (H, W), point_size=point_size)
    meshes = []
    for f in range(N):
        parts = [raster.make_ellipsoid_mesh(m, c, scale_factor, subdivisions, device=dev) for m, c in ellipsoids_per_frame[f]]
        meshes.append(raster.combine_meshes_for_scene(parts))
    fg_depth, fg_mask = raster.render_meshes_depth_batch(meshes, intrinsics, extrinsics, (H, W))
    fg_rgb = torch.zeros_like(bg_rgb)          # the shaded ellipsoid colour is discarded by the reference (:1281)
    _, combined_depth = composite_by_depth(bg_rgb, bg_depth, fg_rgb, fg_depth, fg_mask)
    lo, hi = depth_range if depth_range is not None else compute_global_depth_range([bg_depth, fg_depth, combined_depth])
    out = {
        "background_RGB": bg_rgb,
        "background_depth": visualize_depth_as_grayscale(bg_depth, lo, hi),
        "3D_gaussian_depth": visualize_depth_as_grayscale(fg_depth, lo, hi),
        "merged_mask": merge_bg_and_fg_mask(bg_depth, fg_depth, bg_mask, fg_mask),
        "depth_range": (lo, hi), "bg_depth": bg_depth, "fg_depth": fg_depth, "bg_mask": bg_mask, "fg_mask": fg_mask,
    }
    gp = list(gaussian_params_per_frame) + [{}] * (N - len(gaussian_params_per_frame))
    if any(len(p) for p in gp):
        _, alpha, blend, masked = project_3d_gaussians_to_2d(gp, obj_id_to_color_idx, intrinsics, extrinsics, (W, H),
                                                             threshold=gaussian_mask_threshold, device=dev, background_frames=bg_rgb)
        out["3D_gaussian_RGB"] = masked
        # step 10 (:1353-1373), the preview video: the Gaussian colours alpha-blended over the background
        # (blend_gaussian_projection_with_bg, written by the projection kernel), then depth-composited against the background
        # with the ellipsoid depth, visible where alpha > 0.001 (merge_bg_and_fg_sequences -> composite_by_depth)
        out["gaussian_alpha"], out["gaussian_blend"] = alpha, blend
        out["background_and_3D_gaussian"], _ = composite_by_depth(bg_rgb, bg_depth, blend, fg_depth, alpha > 0.001)
    else:
        out["3D_gaussian_RGB"] = torch.zeros_like(bg_rgb)
        out["background_and_3D_gaussian"] = bg_rgb      # alpha == 0 everywhere: blend == background, nothing visible (:1334-1347)
    return out
