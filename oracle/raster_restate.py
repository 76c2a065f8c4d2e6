"""TEST INFRASTRUCTURE — numpy (fp32, op-by-op) restatement of the point-cloud and mesh rasterisation the reference's
4D control-map renderer delegates to pytorch3d (/root/reference/inference/rendering_4D_control_maps.py:150-338:
PointsRasterizer(radius, points_per_pixel=8) + AlphaCompositor, MeshRasterizer(blur_radius=0, faces_per_pixel=1)).
Only tests/ may import this.

**Parity unpinned.**  pytorch3d is a third-party dependency that is neither vendored in /root/reference nor installed
here (requirements: README.md "pytorch3d"; no pinned version recoverable), so nothing can be run against it.  The
functions below restate its published naive rasterisers (pytorch3d/csrc/rasterize_points/rasterize_points.cu,
rasterize_meshes/rasterize_meshes.cu, compositing/alpha_composite.cu, renderer/points/renderer.py,
renderer/cameras.py PerspectiveCameras with in_ndc=False) and the reference's own camera conversion
(_build_cam_from_extrinsics :338-396).  Brute force over pixels x primitives: small cases only."""
import numpy as np

F = np.float32
K_EPS = F(1e-8)


def pix_to_ndc(i, S1, S2):
    """PixToNonSquareNdc (rasterization_utils.cuh)."""
    rng = F(2.0)
    if S1 > S2:
        rng = F(F(rng * F(S1)) / F(S2))
    off = F(rng / F(2.0))
    return F(-off) + (F(rng) * np.asarray(i, dtype=F) + off) / F(S1)


def camera(K, T, W, H):
    """OpenCV intrinsics (pixels) / world-to-camera -> (R, t, fx, fy, px, py) of the pytorch3d NDC camera
    (_build_cam_from_extrinsics + PerspectiveCameras screen->NDC conversion): view = diag(-1,-1,1)(R x + t)."""
    K, T = np.asarray(K, dtype=F), np.asarray(T, dtype=F)
    s = F(F(min(W, H)) / F(2.0))
    return (T[:3, :3].copy(), T[:3, 3].copy(), F(K[0, 0] / s), F(K[1, 1] / s),
            F(-(K[0, 2] - F(F(W) / F(2.0))) / s), F(-(K[1, 2] - F(F(H) / F(2.0))) / s))


def project(pts, cam):
    """world [N,3] -> (x_ndc, y_ndc, z_view), fp32 without fused multiply-adds, same association as the kernel."""
    R, t, fx, fy, px, py = cam
    p = np.asarray(pts, dtype=F)
    x, y, z = p[:, 0], p[:, 1], p[:, 2]
    xc = ((R[0, 0] * x + R[0, 1] * y) + R[0, 2] * z) + t[0]
    yc = ((R[1, 0] * x + R[1, 1] * y) + R[1, 2] * z) + t[1]
    zc = ((R[2, 0] * x + R[2, 1] * y) + R[2, 2] * z) + t[2]
    with np.errstate(divide="ignore", invalid="ignore"):
        return (fx * -xc) / zc + px, (fy * -yc) / zc + py, zc


def render_points(points, colors, K, T, image_size, radius, background=(0.5, 0.5, 0.5), k_points=8):
    """render_point_cloud_pytorch3d_batch (:243-336) for one frame -> (rgb uint8 [H,W,3], depth f32, mask bool)."""
    H, W = image_size
    radius = F(radius)
    r2 = F(radius * radius)
    px, py, pz = project(points, camera(K, T, W, H))
    xf = pix_to_ndc(W - 1 - np.arange(W), W, H)[None, :, None]
    yf = pix_to_ndc(H - 1 - np.arange(H), H, W)[:, None, None]
    with np.errstate(invalid="ignore", over="ignore"):
        dx, dy = xf - px[None, None, :], yf - py[None, None, :]
        d2 = dx * dx + dy * dy                                              # [H, W, N]
        ok = (d2 < r2) & (pz >= 0)[None, None, :] & np.isfinite(px)[None, None] & np.isfinite(py)[None, None]
    col = np.asarray(colors).astype(F) / F(255.0)
    rgb = np.zeros((H, W, 3), np.uint8)
    depth = np.zeros((H, W), F)
    mask = np.zeros((H, W), bool)
    for yi in range(H):
        for xi in range(W):
            cand = np.nonzero(ok[yi, xi])[0]
            if cand.size == 0:
                rgb[yi, xi] = [np.uint8(min(max(F(F(b) * F(255.0)), 0), 255)) for b in background]
                continue
            order = cand[np.lexsort((cand, pz[cand]))][:k_points]           # nearest z first, index breaks ties
            acc, cum = np.zeros(3, F), F(1.0)
            for n in order:
                alpha = F(F(1.0) - F(d2[yi, xi, n] / r2))
                w = F(cum * alpha)
                acc = (acc + w * col[n]).astype(F)
                cum = F(cum * F(F(1.0) - alpha))
            rgb[yi, xi] = np.clip(acc * F(255.0), 0, 255).astype(np.uint8)
            depth[yi, xi] = pz[order[0]]
            mask[yi, xi] = True
    return rgb, depth, mask


def _edge(px, py, ax, ay, bx, by):
    return (px - ax) * (by - ay) - (py - ay) * (bx - ax)


def render_mesh_depth(verts, faces, K, T, image_size):
    """Depth / mask / pix_to_face of render_meshes_pytorch3d_batch (:150-241) for one frame (the shaded RGB of that
    function feeds nothing the pipeline writes out and is not restated)."""
    H, W = image_size
    vx, vy, vz = project(verts, camera(K, T, W, H))
    faces = np.asarray(faces)
    depth = np.zeros((H, W), F)
    p2f = np.full((H, W), -1, np.int64)
    best = np.full((H, W), np.inf, F)
    px = np.broadcast_to(pix_to_ndc(W - 1 - np.arange(W), W, H)[None, :], (H, W)).astype(F)
    py = np.broadcast_to(pix_to_ndc(H - 1 - np.arange(H), H, W)[:, None], (H, W)).astype(F)
    for f, (i0, i1, i2) in enumerate(faces):
        x0, y0, z0, x1, y1, z1, x2, y2, z2 = vx[i0], vy[i0], vz[i0], vx[i1], vy[i1], vz[i1], vx[i2], vy[i2], vz[i2]
        if not max(z0, z1, z2) >= 0:
            continue
        face_area = _edge(x0, y0, x1, y1, x2, y2)
        if -K_EPS <= face_area <= K_EPS or not np.isfinite(x0 + x1 + x2 + y0 + y1 + y2):
            continue
        area = F(_edge(x2, y2, x0, y0, x1, y1) + K_EPS)
        with np.errstate(invalid="ignore", divide="ignore", over="ignore"):
            b0 = _edge(px, py, x1, y1, x2, y2) / area
            b1 = _edge(px, py, x2, y2, x0, y0) / area
            b2 = _edge(px, py, x0, y0, x1, y1) / area
            w0, w1, w2 = (b0 * z1) * z2, (z0 * b1) * z2, (z0 * z1) * b2
            den = np.maximum((w0 + w1) + w2, K_EPS)
            c0, c1, c2 = w0 / den, w1 / den, w2 / den
            pz = (c0 * z0 + c1 * z1) + c2 * z2
            hit = (pz >= 0) & (c0 > 0) & (c1 > 0) & (c2 > 0) & (pz < best)
        best = np.where(hit, pz, best)
        p2f = np.where(hit, f, p2f)
    mask = p2f >= 0
    depth[mask] = best[mask]
    return depth, mask, p2f


def ico_sphere(level):
    """Unit icosphere (pytorch3d.utils.ico_sphere: 12-vertex icosahedron, `level` midpoint subdivisions, vertices
    re-normalised after every subdivision) -> (verts [V,3] f32, faces [F,3] int64); level 3: 642 / 1280."""
    a, b = 0.5257, 0.8507
    verts = [[-a, b, 0], [a, b, 0], [-a, -b, 0], [a, -b, 0], [0, -a, b], [0, a, b], [0, -a, -b], [0, a, -b],
             [b, 0, -a], [b, 0, a], [-b, 0, -a], [-b, 0, a]]
    faces = [[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6],
             [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10],
             [8, 6, 7], [9, 8, 1]]
    v = np.asarray(verts, dtype=F)
    f = np.asarray(faces, dtype=np.int64)
    for _ in range(level):
        edges = np.sort(np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]]), axis=1)
        uniq, inv = np.unique(edges, axis=0, return_inverse=True)
        mid = ((v[uniq[:, 0]] + v[uniq[:, 1]]) * F(0.5)).astype(F)
        n0 = len(v)
        e01, e12, e20 = (n0 + inv.reshape(3, -1)[i] for i in range(3))
        f = np.concatenate([np.stack([f[:, 0], e01, e20], 1), np.stack([f[:, 1], e12, e01], 1),
                            np.stack([f[:, 2], e20, e12], 1), np.stack([e01, e12, e20], 1)])
        v = np.concatenate([v, mid])
        v = (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(F)
    return v, f


def ellipsoid_mesh(mean, cov, scale_factor=2.0, subdivisions=3):
    """make_ellipsoid_mesh (:62-110): unit icosphere scaled by scale_factor * sqrt(eigenvalues) along the eigenvectors of
    the covariance and moved to the mean."""
    v, f = ico_sphere(subdivisions)
    evals, evecs = np.linalg.eigh(np.asarray(cov, dtype=F))
    axes = F(scale_factor) * np.sqrt(np.clip(evals, 1e-8, None)).astype(F)
    M = (evecs @ np.diag(axes)).astype(F)
    return (v @ M.T + np.asarray(mean, dtype=F)).astype(F), f
