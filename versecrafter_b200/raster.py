"""Point-cloud and ellipsoid-mesh rendering of the reference's 4D control-map renderer on libvc_b200.so kernels
(/root/reference/inference/rendering_4D_control_maps.py).  Function names and argument meaning
follow the reference's `render_point_cloud_pytorch3d_batch` (:243-336), `render_meshes_pytorch3d_batch` (:150-241, depth
and mask outputs), `make_ellipsoid_mesh` (:62-110) and `combine_meshes_for_scene` (:112-148); the rasterisation itself
restates pytorch3d's published algorithm (pytorch3d is absent here: parity unpinned).  No CPU fallback."""
import ctypes

import numpy as np
import torch

from . import _lib, ops


def _f32_host(a, shape):
    a = np.ascontiguousarray(torch.as_tensor(a).detach().cpu().numpy(), dtype=np.float32)
    if a.shape != shape:
        raise _lib.VCError(f"expected shape {shape}, got {a.shape}")
    return a, a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def render_point_cloud_batch(points_3d, colors, Ks, Ts, image_size, point_size=0.01, background_color=(0.5, 0.5, 0.5)):
    """points_3d fp32 [N,3], colors uint8 [N,3] (CUDA), Ks [B,3,3], Ts [B,4,4] -> (rgb uint8 [B,H,W,3], depth fp32
    [B,H,W], mask bool [B,H,W]).  NaN/Inf points are dropped first, as in the reference (:283-288)."""
    H, W = image_size
    if not points_3d.is_cuda or not colors.is_cuda:
        raise _lib.VCError("render_point_cloud_batch: CUDA tensors expected (no CPU fallback)")
    dev = points_3d.device
    pts = points_3d.to(torch.float32)
    valid = torch.isfinite(pts).all(dim=1) & torch.isfinite(colors.float()).all(dim=1)
    pts, col = pts[valid].contiguous(), colors[valid].to(torch.uint8).contiguous()
    B, n = len(Ks), pts.shape[0]
    rgb = torch.empty((B, H, W, 3), dtype=torch.uint8, device=dev)
    depth = torch.empty((B, H, W), dtype=torch.float32, device=dev)
    mask = torch.empty((B, H, W), dtype=torch.uint8, device=dev)
    proj = torch.empty((max(n, 1), 3), dtype=torch.float32, device=dev)
    sel = torch.empty((8, H, W), dtype=torch.int64, device=dev)
    bg = (ctypes.c_float * 3)(*[float(c) for c in background_color])
    lib = _lib.load()
    for b in range(B):
        _, Kp = _f32_host(Ks[b], (3, 3))
        _, Tp = _f32_host(Ts[b], (4, 4))
        _lib.check(lib.vc_render_points(ops._ptr(pts), ops._ptr(col), n, Kp, Tp, W, H, float(point_size), bg, ops._ptr(proj),
                                        ops._ptr(sel), ops._ptr(rgb[b]), ops._ptr(depth[b]), ops._ptr(mask[b]),
                                        ops._stream()), "vc_render_points")
        ops.launch_count += 10 if n else 1
    return rgb, depth, mask.bool()


def render_meshes_depth_batch(meshes_list, Ks, Ts, image_size):
    """meshes_list: per frame None or (verts fp32 [V,3], faces int [F,3]) CUDA tensors -> (depth fp32 [B,H,W], mask bool
    [B,H,W]); frames without a mesh come back empty (:176-181)."""
    H, W = image_size
    B = len(meshes_list)
    dev = next((m[0].device for m in meshes_list if m is not None), torch.device("cuda"))
    depth = torch.zeros((B, H, W), dtype=torch.float32, device=dev)
    mask = torch.zeros((B, H, W), dtype=torch.uint8, device=dev)
    zkey = torch.empty((H, W), dtype=torch.int64, device=dev)
    lib = _lib.load()
    for b, m in enumerate(meshes_list):
        if m is None or m[0].shape[0] == 0:
            continue
        verts = m[0].to(device=dev, dtype=torch.float32).contiguous()
        faces = m[1].to(device=dev, dtype=torch.int32).contiguous()
        vproj = torch.empty_like(verts)
        _, Kp = _f32_host(Ks[b], (3, 3))
        _, Tp = _f32_host(Ts[b], (4, 4))
        _lib.check(lib.vc_render_mesh_depth(ops._ptr(verts), verts.shape[0], ops._ptr(faces), faces.shape[0], Kp, Tp, W, H,
                                            ops._ptr(vproj), ops._ptr(zkey), ops._ptr(depth[b]), ops._ptr(mask[b]), None,
                                            ops._stream()), "vc_render_mesh_depth")
        ops.launch_count += 3
    return depth, mask.bool()


_ICO_CACHE = {}


def ico_sphere(level):
    """Unit icosphere as pytorch3d.utils.ico_sphere builds it (icosahedron + `level` midpoint subdivisions, vertices
    re-normalised each time) -> (verts float32 [V,3], faces int64 [F,3]) numpy arrays."""
    if level in _ICO_CACHE:
        return _ICO_CACHE[level]
    a, b = 0.5257, 0.8507
    v = np.asarray([[-a, b, 0], [a, b, 0], [-a, -b, 0], [a, -b, 0], [0, -a, b], [0, a, b], [0, -a, -b], [0, a, -b],
                    [b, 0, -a], [b, 0, a], [-b, 0, -a], [-b, 0, a]], dtype=np.float32)
    f = np.asarray([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2],
                    [10, 7, 6], [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11],
                    [6, 2, 10], [8, 6, 7], [9, 8, 1]], dtype=np.int64)
    for _ in range(level):
        edges = np.sort(np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]]), axis=1)
        uniq, inv = np.unique(edges, axis=0, return_inverse=True)
        mid = ((v[uniq[:, 0]] + v[uniq[:, 1]]) * np.float32(0.5)).astype(np.float32)
        e01, e12, e20 = (len(v) + inv.reshape(3, -1)[i] for i in range(3))
        f = np.concatenate([np.stack([f[:, 0], e01, e20], 1), np.stack([f[:, 1], e12, e01], 1),
                            np.stack([f[:, 2], e20, e12], 1), np.stack([e01, e12, e20], 1)])
        v = np.concatenate([v, mid])
        v = (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(np.float32)
    _ICO_CACHE[level] = (v, f)
    return v, f


def make_ellipsoid_mesh(mean, cov, scale_factor=2.0, subdivisions=3, device="cuda"):
    """make_ellipsoid_mesh (:62-110) -> (verts fp32 [V,3], faces int32 [F,3]) on `device`."""
    v, f = ico_sphere(subdivisions)
    mean_t = torch.as_tensor(mean, dtype=torch.float32).cpu()
    evals, evecs = torch.linalg.eigh(torch.as_tensor(cov, dtype=torch.float32).cpu())
    axes = scale_factor * torch.sqrt(torch.clamp(evals, min=1e-8))
    M = evecs @ torch.diag(axes)
    verts = torch.from_numpy(v) @ M.T + mean_t
    return verts.to(device), torch.from_numpy(f).to(device=device, dtype=torch.int32)


def combine_meshes_for_scene(mesh_list):
    """combine_meshes_for_scene (:112-148): concatenate vertices, offset face indices."""
    if len(mesh_list) == 0:
        return None
    verts, faces, ofs = [], [], 0
    for v, f in mesh_list:
        verts.append(v)
        faces.append(f + ofs)
        ofs += v.shape[0]
    return torch.cat(verts), torch.cat(faces)
