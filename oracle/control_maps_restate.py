"""TEST INFRASTRUCTURE — torch restatement (CPU, fp32) of the per-pixel parts of the reference's 4D control-map
renderer (/root/reference/inference/rendering_4D_control_maps.py).  Only tests/ may import this.

Pinned: tests/golden/gen_golden_control_maps.py cuts the reference's own functions out of the unmodified source with
`ast` (the module itself cannot be imported: pytorch3d / kornia / matplotlib are absent) and stores their outputs in
tests/golden/control_maps.npz; tests/test_control_maps.py checks every function below against those fixtures.
The tab20 colour table stands in for matplotlib.colormaps['tab20'] (matplotlib is not installed; values recalled
from matplotlib's published table, n/255)."""
import math

import torch

TAB20_255 = [(31, 119, 180), (174, 199, 232), (255, 127, 14), (255, 187, 120), (44, 160, 44), (152, 223, 138),
             (214, 39, 40), (255, 152, 150), (148, 103, 189), (197, 176, 213), (140, 86, 75), (196, 156, 148),
             (227, 119, 194), (247, 182, 210), (127, 127, 127), (199, 199, 199), (188, 189, 34), (219, 219, 141),
             (23, 190, 207), (158, 218, 229)]


def tab20(idx):
    return tuple(c / 255.0 for c in TAB20_255[idx % 20])


def composite_by_depth(bg_rgb, bg_depth, fg_rgb, fg_depth, fg_mask):
    """:438-455 (and the batch variant :398-411)."""
    take_fg = fg_mask & ((bg_depth <= 0) | ((fg_depth > 0) & (fg_depth < bg_depth - 1e-6)))
    out_rgb = bg_rgb.clone()
    out_rgb[take_fg] = fg_rgb[take_fg]
    out_depth = bg_depth.clone()
    out_depth[take_fg] = fg_depth[take_fg]
    return out_rgb, out_depth


def merge_bg_and_fg_mask(bg_depth, fg_depth, bg_mask, fg_mask):
    """:736-763 for one frame -> uint8 [H, W, 3]."""
    take_fg = fg_mask & ((bg_depth <= 0) | ((fg_depth > 0) & (fg_depth < bg_depth - 1e-6)))
    out = ~bg_mask
    out = out.clone()
    out[take_fg] = fg_mask[take_fg]
    return (torch.stack([out, out, out], dim=-1) * 255).to(torch.uint8)


def depth_to_gray(depth, min_depth, max_depth):
    """visualize_depth_as_grayscale (:487-539) for one frame with a given (python float) range."""
    disp = torch.where(depth > 0, 1.0 / depth, torch.tensor(0.0))
    if max_depth > 0 and min_depth > 0:
        min_disp = 1.0 / max_depth
        max_disp = 1.0 / min_depth
        disp = (disp - min_disp) / (max_disp - min_disp + 1e-8)
    disp = torch.clamp(disp, 0, 1)
    g = (disp * 255).to(torch.uint8)
    return g.unsqueeze(-1).repeat(1, 1, 3)


def gaussian_2d_params(mean, cov, K, R, t):
    """Projection of one 3D Gaussian (compute_probability_density_map_gpu :765-852): returns None when the reference
    skips it (z <= 0.5 or det > 1e11), else (mean_2d [2], cov_inv [2,2], coeff, depth z) as fp32 tensors / floats."""
    t_vec = t.squeeze() if t.dim() == 2 else t
    m_cam = R @ mean + t_vec
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
    cov_inv = torch.linalg.inv(cov_2d)
    coeff = 1.0 / (2 * torch.pi * torch.sqrt(det))
    return mean_2d, cov_inv, coeff, float(z)


def density_map(params, W, H):
    """pdf of one projected Gaussian on the pixel grid (:854-874) -> [H, W]; params=None -> zeros."""
    if params is None:
        return torch.zeros(H, W)
    mean_2d, cov_inv, coeff, _ = params
    u, v = torch.meshgrid(torch.arange(W, dtype=torch.float32), torch.arange(H, dtype=torch.float32), indexing="xy")
    diff = (torch.stack([u, v], dim=-1) - mean_2d).reshape(-1, 2)
    mahal = torch.sum((diff @ cov_inv) * diff, dim=1)
    pdf = (coeff * torch.exp(-0.5 * mahal)).reshape(H, W)
    return torch.nan_to_num(pdf, nan=0.0, posinf=0.0, neginf=0.0)


def project_frame(gaussians, colors, K, T, W, H, threshold):
    """One frame of project_3d_gaussians_to_2d (:603-692).  gaussians: list of (mean [3], cov [3,3]) in dict order,
    colors: list of (r,g,b) floats.  Returns (rgb uint8 [H,W,3], alpha fp32 [H,W])."""
    R, t = T[:3, :3], T[:3, 3:4]
    items = []
    for (mean, cov), col in zip(gaussians, colors):
        depth_z = float((R @ mean + t.squeeze())[2])
        if depth_z > 0:
            dm = density_map(gaussian_2d_params(mean, cov, K, R, t), W, H)
            mx = dm.max()
            items.append((dm / (mx + 1e-8) if mx > 0 else dm, torch.tensor(col, dtype=torch.float32), depth_z))
    items.sort(key=lambda it: it[2], reverse=True)
    rgb = torch.zeros(H, W, 3)
    alpha = torch.zeros(H, W)
    for dn, col, _ in items:
        a = torch.where(dn > threshold, (dn - threshold) / (1.0 - threshold + 1e-8), torch.zeros_like(dn)).clamp(0.0, 1.0)
        ae = a.unsqueeze(-1)
        rgb = col.view(1, 1, 3) * ae + rgb * (1 - ae)
        alpha = a + alpha * (1 - a)
    return (rgb.clamp(0, 1) * 255).to(torch.uint8), alpha.clamp(0, 1)


def blend_with_bg(g_rgb, alpha, bg_rgb):
    """blend_gaussian_projection_with_bg (:697-733) for one frame."""
    a = alpha.unsqueeze(-1)
    m = (g_rgb.float() / 255.0) * a + (bg_rgb.float() / 255.0) * (1 - a)
    return (m.clamp(0, 1) * 255).to(torch.uint8)


def masked_projection(g_rgb, alpha):
    """main(): (rgb/255 * alpha * 255).to(uint8) — the frames written to 3D_gaussian_RGB.mp4."""
    return ((g_rgb.float() / 255.0) * alpha.unsqueeze(-1) * 255).to(torch.uint8)
