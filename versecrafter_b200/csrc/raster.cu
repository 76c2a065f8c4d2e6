// Point-cloud splatting and triangle-mesh depth rasterisation of the 4D control-map renderer as CUDA kernels
// (/root/reference/inference/rendering_4D_control_maps.py:150-338).  The reference calls
// pytorch3d (PointsRasterizer + AlphaCompositor, MeshRasterizer); pytorch3d is not in the reference tree and not
// installed here, so these kernels follow its PUBLISHED algorithm (naive rasterisers rasterize_points.cu /
// rasterize_meshes.cu, compositing/alpha_composite.cu) — parity unpinned, see oracle/raster_restate.py.
//
// Conventions (pytorch3d): NDC +X left, +Y up; the short image side spans [-1, 1]; pixel (yi, xi) has the centre
//   xf = pix_to_ndc(W-1-xi, W, H), yf = pix_to_ndc(H-1-yi, H, W);  depth = view-space z.
// Camera of _build_cam_from_extrinsics (:338-396): view = diag(-1,-1,1) * (R x + t) for the OpenCV world-to-camera
// [R|t]; screen-space intrinsics are converted to NDC with s = min(H, W) / 2.
//
// Points: (1) project every point once per frame; (2) K = 8 selection passes: pass k keeps, per pixel, the smallest
// (z, index) key larger than the key chosen in pass k-1 (64-bit atomicMin) — the K nearest points in z order without
// per-pixel lists or sorting; (3) per-pixel front-to-back alpha compositing with weights 1 - d^2 / r^2.
// Meshes: one thread per face walks the face's pixel bounding box, perspective-correct barycentrics, 64-bit atomicMin
// of (z, face) per covered pixel; a resolve pass writes depth and mask.
#include "common.cuh"

namespace {

constexpr int K_POINTS = 8;
constexpr unsigned long long EMPTY_KEY = 0xFFFFFFFFFFFFFFFFull;
constexpr float kEps = 1e-8f;

struct Cam {
  float r[9];   // OpenCV world-to-camera rotation, row-major
  float t[3];
  float fx, fy, px, py;  // NDC focal length / principal point (already converted)
};

__device__ __forceinline__ float pix_to_ndc(int i, int S1, int S2) {
  float range = 2.0f;
  if (S1 > S2) range = __fdiv_rn(__fmul_rn(range, (float)S1), (float)S2);
  const float offset = __fdiv_rn(range, 2.0f);
  return __fadd_rn(-offset, __fdiv_rn(__fadd_rn(__fmul_rn(range, (float)i), offset), (float)S1));
}

// world point -> (x_ndc, y_ndc, z_view), evaluated without fma contraction so the CPU restatement can match bit-wise
__device__ __forceinline__ float3 project(const Cam& c, float x, float y, float z) {
  const float xc = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(c.r[0], x), __fmul_rn(c.r[1], y)), __fmul_rn(c.r[2], z)), c.t[0]);
  const float yc = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(c.r[3], x), __fmul_rn(c.r[4], y)), __fmul_rn(c.r[5], z)), c.t[1]);
  const float zc = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(c.r[6], x), __fmul_rn(c.r[7], y)), __fmul_rn(c.r[8], z)), c.t[2]);
  float3 o;
  o.x = __fadd_rn(__fdiv_rn(__fmul_rn(c.fx, -xc), zc), c.px);
  o.y = __fadd_rn(__fdiv_rn(__fmul_rn(c.fy, -yc), zc), c.py);
  o.z = zc;
  return o;
}

__global__ void __launch_bounds__(256) project_kernel(const float* __restrict__ pts, long long n, const __grid_constant__ Cam cam,
                                                     float* __restrict__ out /* [n,3] */) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float3 q = project(cam, pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]);
    out[3 * i] = q.x;
    out[3 * i + 1] = q.y;
    out[3 * i + 2] = q.z;
  }
}

__device__ __forceinline__ float dist2_ndc(float px, float py, int xi, int yi, int W, int H) {
  const float dx = __fsub_rn(pix_to_ndc(W - 1 - xi, W, H), px);
  const float dy = __fsub_rn(pix_to_ndc(H - 1 - yi, H, W), py);
  return __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
}

// conservative pixel range [lo, hi] whose centres can lie within `rad` (NDC) of NDC coordinate c along an axis of S1
// pixels (other axis S2): centre(i) = off - (range/S1) * (i + 0.5)
__device__ __forceinline__ void pixel_span(float c, float rad, int S1, int S2, int& lo, int& hi) {
  const float range = S1 > S2 ? 2.0f * (float)S1 / (float)S2 : 2.0f;
  const float step = range / (float)S1;
  const float u = (0.5f * range - c) / step - 0.5f;   // fractional pixel index of c
  const float rp = rad / step + 1.0f;
  lo = (int)floorf(fminf(fmaxf(u - rp, 0.0f), (float)S1));        // clamped before the conversion: far-off points
  hi = (int)ceilf(fminf(fmaxf(u + rp, -1.0f), (float)(S1 - 1)));  // give an empty span, never an overflow
}

// selection pass: sel_prev == nullptr for the first pass
__global__ void __launch_bounds__(256) points_select_kernel(const float* __restrict__ proj, long long n, int W, int H,
                                                           float radius, const unsigned long long* __restrict__ sel_prev,
                                                           unsigned long long* __restrict__ sel_cur) {
  const float r2 = __fmul_rn(radius, radius);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float px = proj[3 * i], py = proj[3 * i + 1], pz = proj[3 * i + 2];
    if (!(pz >= 0.0f) || !isfinite(px) || !isfinite(py)) continue;   // behind the camera (pz < 0) or degenerate
    int x0, x1, y0, y1;
    pixel_span(px, radius, W, H, x0, x1);
    pixel_span(py, radius, H, W, y0, y1);
    const unsigned long long key = ((unsigned long long)__float_as_uint(pz) << 32) | (unsigned long long)(i + 1);
    for (int yi = y0; yi <= y1; ++yi) {
      for (int xi = x0; xi <= x1; ++xi) {
        if (dist2_ndc(px, py, xi, yi, W, H) < r2) {
          const long long pix = (long long)yi * W + xi;
          if (sel_prev == nullptr || key > sel_prev[pix]) atomicMin(&sel_cur[pix], key);
        }
      }
    }
  }
}

// alpha_composite over the K selected points (front to back) + background fill + uint8 quantisation (:318-327)
__global__ void __launch_bounds__(256) points_composite_kernel(const float* __restrict__ proj, const uint8_t* __restrict__ colors,
                                                              const unsigned long long* __restrict__ sel /* [K, H*W] */,
                                                              int W, int H, float radius, float bg_r, float bg_g, float bg_b,
                                                              uint8_t* __restrict__ rgb, float* __restrict__ depth,
                                                              uint8_t* __restrict__ mask) {
  const long long n_pix = (long long)W * H;
  const float r2 = __fmul_rn(radius, radius);
  for (long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x; pix < n_pix; pix += (long long)gridDim.x * blockDim.x) {
    const int yi = (int)(pix / W), xi = (int)(pix % W);
    float acc[3] = {0.f, 0.f, 0.f};
    float cum = 1.0f;
    float z0 = 0.0f;
    bool any = false;
#pragma unroll 1
    for (int k = 0; k < K_POINTS; ++k) {
      const unsigned long long key = sel[(long long)k * n_pix + pix];
      if (key == EMPTY_KEY) break;
      const long long idx = (long long)(key & 0xFFFFFFFFull) - 1;
      if (!any) {
        any = true;
        z0 = __uint_as_float((unsigned int)(key >> 32));
      }
      const float d2 = dist2_ndc(proj[3 * idx], proj[3 * idx + 1], xi, yi, W, H);
      const float alpha = __fsub_rn(1.0f, __fdiv_rn(d2, r2));
      const float wgt = __fmul_rn(cum, alpha);
#pragma unroll
      for (int c = 0; c < 3; ++c)
        acc[c] = __fadd_rn(acc[c], __fmul_rn(wgt, __fdiv_rn((float)colors[3 * idx + c], 255.0f)));
      cum = __fmul_rn(cum, __fsub_rn(1.0f, alpha));
    }
    if (!any) {
      acc[0] = bg_r; acc[1] = bg_g; acc[2] = bg_b;
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) rgb[3 * pix + c] = (uint8_t)fminf(fmaxf(__fmul_rn(acc[c], 255.0f), 0.0f), 255.0f);
    depth[pix] = any ? z0 : 0.0f;
    mask[pix] = any ? 1 : 0;
  }
}


// ---------------------------------------------------------------------------------------------------------------
// Triangle meshes (MeshRasterizer, blur_radius = 0, faces_per_pixel = 1, perspective-correct, no culling/clipping)
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float edge_fn(float px, float py, float ax, float ay, float bx, float by) {
  return __fsub_rn(__fmul_rn(__fsub_rn(px, ax), __fsub_rn(by, ay)), __fmul_rn(__fsub_rn(py, ay), __fsub_rn(bx, ax)));
}

__global__ void __launch_bounds__(128) mesh_raster_kernel(const float* __restrict__ vproj, const int* __restrict__ faces,
                                                         int n_faces, int W, int H, unsigned long long* __restrict__ zkey) {
  for (int f = blockIdx.x * blockDim.x + threadIdx.x; f < n_faces; f += gridDim.x * blockDim.x) {
    const int i0 = faces[3 * f], i1 = faces[3 * f + 1], i2 = faces[3 * f + 2];
    const float x0 = vproj[3 * i0], y0 = vproj[3 * i0 + 1], z0 = vproj[3 * i0 + 2];
    const float x1 = vproj[3 * i1], y1 = vproj[3 * i1 + 1], z1 = vproj[3 * i1 + 2];
    const float x2 = vproj[3 * i2], y2 = vproj[3 * i2 + 1], z2 = vproj[3 * i2 + 2];
    if (!(fmaxf(z0, fmaxf(z1, z2)) >= 0.0f)) continue;                       // zmax < 0
    const float face_area = edge_fn(x0, y0, x1, y1, x2, y2);                 // EdgeFunctionForward(v0, v1, v2)
    if (face_area <= kEps && face_area >= -kEps) continue;                   // zero-area face
    if (!isfinite(x0 + x1 + x2 + y0 + y1 + y2)) continue;
    int xa, xb, ya, yb, t0, t1;
    pixel_span(fmaxf(x0, fmaxf(x1, x2)), 0.0f, W, H, xa, t1);                // NDC +x is image-left: max x -> min column
    pixel_span(fminf(x0, fminf(x1, x2)), 0.0f, W, H, t0, xb);
    pixel_span(fmaxf(y0, fmaxf(y1, y2)), 0.0f, H, W, ya, t1);
    pixel_span(fminf(y0, fminf(y1, y2)), 0.0f, H, W, t0, yb);
    const float area = __fadd_rn(edge_fn(x2, y2, x0, y0, x1, y1), kEps);     // BarycentricCoordsForward
    for (int yi = ya; yi <= yb; ++yi) {
      const float py = pix_to_ndc(H - 1 - yi, H, W);
      for (int xi = xa; xi <= xb; ++xi) {
        const float px = pix_to_ndc(W - 1 - xi, W, H);
        const float b0 = __fdiv_rn(edge_fn(px, py, x1, y1, x2, y2), area);
        const float b1 = __fdiv_rn(edge_fn(px, py, x2, y2, x0, y0), area);
        const float b2 = __fdiv_rn(edge_fn(px, py, x0, y0, x1, y1), area);
        const float w0t = __fmul_rn(__fmul_rn(b0, z1), z2);
        const float w1t = __fmul_rn(__fmul_rn(z0, b1), z2);
        const float w2t = __fmul_rn(__fmul_rn(z0, z1), b2);
        const float den = fmaxf(__fadd_rn(__fadd_rn(w0t, w1t), w2t), kEps);
        const float c0 = __fdiv_rn(w0t, den), c1 = __fdiv_rn(w1t, den), c2 = __fdiv_rn(w2t, den);
        const float pz = __fadd_rn(__fadd_rn(__fmul_rn(c0, z0), __fmul_rn(c1, z1)), __fmul_rn(c2, z2));
        if (!(pz >= 0.0f)) continue;
        if (!(c0 > 0.0f && c1 > 0.0f && c2 > 0.0f)) continue;                // blur_radius = 0: strictly inside only
        atomicMin(&zkey[(long long)yi * W + xi], ((unsigned long long)__float_as_uint(pz) << 32) | (unsigned)(f + 1));
      }
    }
  }
}

__global__ void __launch_bounds__(256) mesh_resolve_kernel(const unsigned long long* __restrict__ zkey, long long n_pix,
                                                          float* __restrict__ depth, uint8_t* __restrict__ mask,
                                                          int* __restrict__ pix_to_face) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_pix; i += (long long)gridDim.x * blockDim.x) {
    const unsigned long long k = zkey[i];
    const bool hit = k != EMPTY_KEY;
    depth[i] = hit ? __uint_as_float((unsigned int)(k >> 32)) : 0.0f;        // depth[~mask] = 0 (:223-224)
    mask[i] = hit ? 1 : 0;
    if (pix_to_face) pix_to_face[i] = hit ? (int)(k & 0xFFFFFFFFull) - 1 : -1;
  }
}

int grid_for(long long items) {
  long long blocks = (items + 255) / 256;
  if (blocks > 148LL * 16) blocks = 148LL * 16;
  return (int)(blocks < 1 ? 1 : blocks);
}

Cam make_cam(const float* K, const float* T, int W, int H) {
  Cam c;
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) c.r[3 * i + j] = T[4 * i + j];
    c.t[i] = T[4 * i + 3];
  }
  const float s = (float)(W < H ? W : H) / 2.0f;
  c.fx = K[0] / s;
  c.fy = K[4] / s;
  c.px = -(K[2] - (float)W / 2.0f) / s;
  c.py = -(K[5] - (float)H / 2.0f) / s;
  return c;
}

}  // namespace

extern "C" int vc_render_points(const void* points, const void* colors, long long n_points, const float* K, const float* T,
                                int W, int H, float radius, const float* background, void* proj_scratch,
                                void* sel_scratch, void* out_rgb, void* out_depth, void* out_mask, void* stream) {
  VC_CHECK_ARG(K && T && background && sel_scratch && out_rgb && out_depth && out_mask, "vc_render_points: null operand");
  VC_CHECK_ARG(n_points <= 0 || (points && colors && proj_scratch), "vc_render_points: null point cloud operand");
  VC_CHECK_ARG(n_points >= 0 && n_points < 0xFFFFFFFFll && W > 0 && H > 0 && radius > 0.0f,
               "vc_render_points: bad sizes n=%lld W=%d H=%d radius=%g", n_points, W, H, (double)radius);
  cudaStream_t s = (cudaStream_t)stream;
  const Cam cam = make_cam(K, T, W, H);
  const long long n_pix = (long long)W * H;
  unsigned long long* sel = (unsigned long long*)sel_scratch;
  cudaError_t e = cudaMemsetAsync(sel, 0xFF, (size_t)K_POINTS * n_pix * sizeof(unsigned long long), s);
  if (e != cudaSuccess) {
    vc_host::set_error("vc_render_points: memset failed: %s", cudaGetErrorString(e));
    return (int)e;
  }
  if (n_points > 0) {
    project_kernel<<<grid_for(n_points), 256, 0, s>>>((const float*)points, n_points, cam, (float*)proj_scratch);
    for (int k = 0; k < K_POINTS; ++k)
      points_select_kernel<<<grid_for(n_points), 256, 0, s>>>((const float*)proj_scratch, n_points, W, H, radius,
                                                              k == 0 ? nullptr : sel + (long long)(k - 1) * n_pix,
                                                              sel + (long long)k * n_pix);
  }
  points_composite_kernel<<<grid_for(n_pix), 256, 0, s>>>((const float*)proj_scratch, (const uint8_t*)colors, sel, W, H, radius,
                                                          background[0], background[1], background[2], (uint8_t*)out_rgb,
                                                          (float*)out_depth, (uint8_t*)out_mask);
  VC_RETURN_LAUNCH_STATUS("vc_render_points");
}

extern "C" int vc_render_mesh_depth(const void* verts, int n_verts, const void* faces, int n_faces, const float* K,
                                    const float* T, int W, int H, void* vproj_scratch, void* zkey_scratch,
                                    void* out_depth, void* out_mask, void* out_pix_to_face, void* stream) {
  VC_CHECK_ARG(K && T && zkey_scratch && out_depth && out_mask && W > 0 && H > 0, "vc_render_mesh_depth: bad operand");
  VC_CHECK_ARG(n_verts >= 0 && n_faces >= 0 && (n_faces == 0 || (verts && faces && vproj_scratch)),
               "vc_render_mesh_depth: bad mesh (%d verts, %d faces)", n_verts, n_faces);
  cudaStream_t s = (cudaStream_t)stream;
  const Cam cam = make_cam(K, T, W, H);
  const long long n_pix = (long long)W * H;
  cudaError_t e = cudaMemsetAsync(zkey_scratch, 0xFF, (size_t)n_pix * sizeof(unsigned long long), s);
  if (e != cudaSuccess) {
    vc_host::set_error("vc_render_mesh_depth: memset failed: %s", cudaGetErrorString(e));
    return (int)e;
  }
  if (n_faces > 0) {
    project_kernel<<<grid_for(n_verts), 256, 0, s>>>((const float*)verts, n_verts, cam, (float*)vproj_scratch);
    mesh_raster_kernel<<<(n_faces + 127) / 128, 128, 0, s>>>((const float*)vproj_scratch, (const int*)faces, n_faces, W, H,
                                                             (unsigned long long*)zkey_scratch);
  }
  mesh_resolve_kernel<<<grid_for(n_pix), 256, 0, s>>>((const unsigned long long*)zkey_scratch, n_pix, (float*)out_depth,
                                                      (uint8_t*)out_mask, (int*)out_pix_to_face);
  VC_RETURN_LAUNCH_STATUS("vc_render_mesh_depth");
}
