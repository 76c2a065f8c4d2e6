"""TEST INFRASTRUCTURE — full-resolution variants of oracle/raster_restate.py's brute-force rasterisers: the same
arithmetic (same `project`, `pix_to_ndc`, same tests per pixel), evaluated only inside each primitive's pixel bounding box,
so a 1280x720 frame with 10^6 points / 5 k triangles finishes in seconds on the CPU.  tests/test_demo_scene.py checks them
against the brute-force functions on small cases and uses them to pin the renderer chain against the control maps the
reference itself rendered for its demo scene (tests/golden/demo_scene/reference_maps/*.mp4).  Depth and mask only (the
demo outputs hold no background RGB).  Only tests/ may import this."""
import numpy as np

from . import raster_restate as rr

F = np.float32


def _pixel_of_ndc(v, S1, S2):
    """Inverse of pix_to_ndc up to rounding: fractional pixel index i with pix_to_ndc(i) == v."""
    rng = 2.0 * S1 / S2 if S1 > S2 else 2.0
    off = rng / 2.0
    return ((np.asarray(v, dtype=np.float64) + off) * S1 - off) / rng


def render_points_depth(points, K, T, image_size, radius):
    """(depth f32 [H,W], mask bool) of rr.render_points: per pixel the nearest z among points with d^2 < r^2, z >= 0."""
    H, W = image_size
    radius = F(radius)
    r2 = F(radius * radius)
    px, py, pz = rr.project(points, rr.camera(K, T, W, H))
    ok = np.isfinite(px) & np.isfinite(py) & (pz >= 0)
    px, py, pz = px[ok], py[ok], pz[ok]
    xs = rr.pix_to_ndc(W - 1 - np.arange(W), W, H)   # NDC x of pixel column xi
    ys = rr.pix_to_ndc(H - 1 - np.arange(H), H, W)
    # fractional pixel of every point (columns run against NDC x)
    cx = W - 1 - _pixel_of_ndc(px, W, H)
    cy = H - 1 - _pixel_of_ndc(py, H, W)
    rad_px = float(radius) * min(W, H) / 2.0
    reach = int(np.ceil(rad_px)) + 1
    best = np.full(H * W, np.inf, F)
    bx, by = np.floor(cx).astype(np.int64), np.floor(cy).astype(np.int64)
    for dy in range(-reach, reach + 2):
        yi = by + dy
        vy = (yi >= 0) & (yi < H)
        for dx in range(-reach, reach + 2):
            xi = bx + dx
            v = vy & (xi >= 0) & (xi < W)
            if not v.any():
                continue
            xv, yv = xi[v], yi[v]
            ddx, ddy = xs[xv] - px[v], ys[yv] - py[v]
            hit = (ddx * ddx + ddy * ddy) < r2
            np.minimum.at(best, (yv * W + xv)[hit], pz[v][hit])
    best = best.reshape(H, W)
    mask = np.isfinite(best)
    return np.where(mask, best, F(0)).astype(F), mask


def render_mesh_depth(verts, faces, K, T, image_size):
    """(depth, mask) of rr.render_mesh_depth, each face evaluated on its bounding box only."""
    H, W = image_size
    vx, vy, vz = rr.project(verts, rr.camera(K, T, W, H))
    faces = np.asarray(faces)
    best = np.full((H, W), np.inf, F)
    pxs = rr.pix_to_ndc(W - 1 - np.arange(W), W, H).astype(F)
    pys = rr.pix_to_ndc(H - 1 - np.arange(H), H, W).astype(F)
    for i0, i1, i2 in faces:
        x0, y0, z0, x1, y1, z1, x2, y2, z2 = vx[i0], vy[i0], vz[i0], vx[i1], vy[i1], vz[i1], vx[i2], vy[i2], vz[i2]
        if not max(z0, z1, z2) >= 0:
            continue
        face_area = rr._edge(x0, y0, x1, y1, x2, y2)
        if -rr.K_EPS <= face_area <= rr.K_EPS or not np.isfinite(x0 + x1 + x2 + y0 + y1 + y2):
            continue
        fx = W - 1 - _pixel_of_ndc(np.array([x0, x1, x2]), W, H)
        fy = H - 1 - _pixel_of_ndc(np.array([y0, y1, y2]), H, W)
        xa, xb = max(int(np.floor(fx.min())) - 1, 0), min(int(np.ceil(fx.max())) + 2, W)
        ya, yb = max(int(np.floor(fy.min())) - 1, 0), min(int(np.ceil(fy.max())) + 2, H)
        if xa >= xb or ya >= yb:
            continue
        px, py = np.meshgrid(pxs[xa:xb], pys[ya:yb])
        area = F(rr._edge(x2, y2, x0, y0, x1, y1) + rr.K_EPS)
        with np.errstate(invalid="ignore", divide="ignore", over="ignore"):
            b0 = rr._edge(px, py, x1, y1, x2, y2) / area
            b1 = rr._edge(px, py, x2, y2, x0, y0) / area
            b2 = rr._edge(px, py, x0, y0, x1, y1) / area
            w0, w1, w2 = (b0 * z1) * z2, (z0 * b1) * z2, (z0 * z1) * b2
            den = np.maximum((w0 + w1) + w2, rr.K_EPS)
            c0, c1, c2 = w0 / den, w1 / den, w2 / den
            pz = (c0 * z0 + c1 * z1) + c2 * z2
            sub = best[ya:yb, xa:xb]
            hit = (pz >= 0) & (c0 > 0) & (c1 > 0) & (c2 > 0) & (pz < sub)
        best[ya:yb, xa:xb] = np.where(hit, pz, sub)
    mask = np.isfinite(best)
    return np.where(mask, best, F(0)).astype(F), mask
