// Per-pixel control-map kernels of the 4D control-map renderer (reference:
// /root/reference/inference/rendering_4D_control_maps.py).  These are the parts of that file that are plain tensor
// arithmetic in the reference (no pytorch3d): depth-tested background/foreground compositing and the merged mask,
// the depth -> inverse-depth gray map, and the 3D-Gaussian -> 2D density projection with far-to-near alpha compositing
// and the blend over the background.  All HBM-bound, one thread per pixel, coalesced.
// Compiled WITHOUT --use_fast_math (build.py): the reference's fp32 divisions and exp are IEEE / libm-accurate.
#include "common.cuh"

namespace {

// composite_by_depth (:438-455) + merge_bg_and_fg_mask (:736-763) in one pass over the frames.
//   take_fg = fg_mask & ((bg_depth <= 0) | ((fg_depth > 0) & (fg_depth < bg_depth - 1e-6)))
//   out_rgb/out_depth = take_fg ? fg : bg ;  out_mask = take_fg ? fg_mask : ~bg_mask  -> x255, 3 channels
__global__ void __launch_bounds__(256) composite_kernel(const uint8_t* __restrict__ bg_rgb, const float* __restrict__ bg_depth,
                                                       const uint8_t* __restrict__ bg_mask, const uint8_t* __restrict__ fg_rgb,
                                                       const float* __restrict__ fg_depth, const uint8_t* __restrict__ fg_mask,
                                                       uint8_t* __restrict__ out_rgb, float* __restrict__ out_depth,
                                                       uint8_t* __restrict__ out_mask_rgb, long long n_pix) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_pix; i += (long long)gridDim.x * blockDim.x) {
    const float bd = bg_depth[i], fd = fg_depth[i];
    const bool fm = fg_mask[i] != 0;
    const bool take_fg = fm && ((bd <= 0.0f) || ((fd > 0.0f) && (fd < bd - 1e-6f)));
    if (out_rgb) {
      const uint8_t* src = (take_fg ? fg_rgb : bg_rgb) + 3 * i;
      out_rgb[3 * i + 0] = src[0];
      out_rgb[3 * i + 1] = src[1];
      out_rgb[3 * i + 2] = src[2];
    }
    if (out_depth) out_depth[i] = take_fg ? fd : bd;
    if (out_mask_rgb) {
      const uint8_t m = (take_fg ? fm : (bg_mask[i] == 0)) ? 255 : 0;
      out_mask_rgb[3 * i + 0] = m;
      out_mask_rgb[3 * i + 1] = m;
      out_mask_rgb[3 * i + 2] = m;
    }
  }
}

// visualize_depth_as_grayscale (:487-539) with a given global range:
//   disp = depth > 0 ? 1/depth : 0;  if (max > 0 && min > 0) disp = (disp - 1/max) / (1/min - 1/max + 1e-8)
//   gray = uint8(clamp(disp, 0, 1) * 255), replicated to 3 channels.  min_disp / denom are computed by the caller in
//   double (python floats in the reference) and rounded to fp32 as torch does for python-scalar operands.
__global__ void __launch_bounds__(256) depth_gray_kernel(const float* __restrict__ depth, uint8_t* __restrict__ out,
                                                        long long n_pix, float min_disp, float denom, int normalise) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_pix; i += (long long)gridDim.x * blockDim.x) {
    const float d = depth[i];
    float v = d > 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    if (normalise) v = __fdiv_rn(__fsub_rn(v, min_disp), denom);
    v = fminf(fmaxf(v, 0.0f), 1.0f);
    const uint8_t g = (uint8_t)(__fmul_rn(v, 255.0f));
    out[3 * i + 0] = g;
    out[3 * i + 1] = g;
    out[3 * i + 2] = g;
  }
}

constexpr int MAX_G = 32;  // Gaussians per frame
struct GaussFrame {
  int n;                  // Gaussians of this frame, already sorted far -> near (:668)
  float mx[MAX_G], my[MAX_G];              // projected mean (pixels)
  float i00[MAX_G], i01[MAX_G], i10[MAX_G], i11[MAX_G];  // inverse 2-D covariance
  float coeff[MAX_G];                      // 1 / (2 pi sqrt(det))
  float r[MAX_G], g[MAX_G], b[MAX_G];      // colour in [0,1]
};

__device__ __forceinline__ float gauss_pdf(const GaussFrame& f, int k, float u, float v) {
  const float d0 = u - f.mx[k], d1 = v - f.my[k];
  // (diff @ cov_inv) * diff summed over the 2 columns (:866), no FMA contraction (torch evaluates op by op)
  const float t0 = __fadd_rn(__fmul_rn(d0, f.i00[k]), __fmul_rn(d1, f.i10[k]));
  const float t1 = __fadd_rn(__fmul_rn(d0, f.i01[k]), __fmul_rn(d1, f.i11[k]));
  const float m = __fadd_rn(__fmul_rn(t0, d0), __fmul_rn(t1, d1));
  float p = __fmul_rn(f.coeff[k], expf(__fmul_rn(-0.5f, m)));
  if (!isfinite(p)) p = 0.0f;  // torch.nan_to_num(nan=0, posinf=0, neginf=0) (:795)
  return p;
}

// pass 1: per-Gaussian maximum of the density over the image (:655); non-negative floats order like their bits
__global__ void __launch_bounds__(256) gauss_max_kernel(const __grid_constant__ GaussFrame f, int W, int H,
                                                       unsigned int* __restrict__ max_bits) {
  __shared__ unsigned int smax[MAX_G];
  if (threadIdx.x < MAX_G) smax[threadIdx.x] = 0u;
  __syncthreads();
  const long long n_pix = (long long)W * H;
  float best[MAX_G];
  for (int k = 0; k < f.n; ++k) best[k] = 0.0f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_pix; i += (long long)gridDim.x * blockDim.x) {
    const float u = (float)(i % W), v = (float)(i / W);
    for (int k = 0; k < f.n; ++k) best[k] = fmaxf(best[k], gauss_pdf(f, k, u, v));
  }
  for (int k = 0; k < f.n; ++k) {
    float b = best[k];
    for (int o = 16; o > 0; o >>= 1) b = fmaxf(b, __shfl_xor_sync(0xffffffffu, b, o));
    if ((threadIdx.x & 31) == 0) atomicMax(&smax[k], __float_as_uint(b));
  }
  __syncthreads();
  if (threadIdx.x < f.n) atomicMax(&max_bits[threadIdx.x], smax[threadIdx.x]);
}

// pass 2: normalise, threshold to alpha, composite far -> near (:657-690), quantise, blend over the background
// (blend_gaussian_projection_with_bg :697-733) and the masked projection frame (main(): rgb/255 * alpha * 255).
__global__ void __launch_bounds__(256) gauss_composite_kernel(const __grid_constant__ GaussFrame f, int W, int H,
                                                             const unsigned int* __restrict__ max_bits, float threshold,
                                                             const uint8_t* __restrict__ bg_rgb,
                                                             uint8_t* __restrict__ out_rgb, float* __restrict__ out_alpha,
                                                             uint8_t* __restrict__ out_blend,
                                                             uint8_t* __restrict__ out_masked) {
  const long long n_pix = (long long)W * H;
  const float inv_den = __fadd_rn(__fsub_rn(1.0f, threshold), 1e-8f);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_pix; i += (long long)gridDim.x * blockDim.x) {
    const float u = (float)(i % W), v = (float)(i / W);
    float cr = 0.0f, cg = 0.0f, cb = 0.0f, alpha = 0.0f;
    for (int k = 0; k < f.n; ++k) {
      const float mx = __uint_as_float(max_bits[k]);
      float dn = gauss_pdf(f, k, u, v);
      if (mx > 0.0f) dn = __fdiv_rn(dn, __fadd_rn(mx, 1e-8f));
      float a = dn > threshold ? __fdiv_rn(__fsub_rn(dn, threshold), inv_den) : 0.0f;
      a = fminf(fmaxf(a, 0.0f), 1.0f);
      const float ia = __fsub_rn(1.0f, a);
      cr = __fadd_rn(__fmul_rn(f.r[k], a), __fmul_rn(cr, ia));
      cg = __fadd_rn(__fmul_rn(f.g[k], a), __fmul_rn(cg, ia));
      cb = __fadd_rn(__fmul_rn(f.b[k], a), __fmul_rn(cb, ia));
      alpha = __fadd_rn(a, __fmul_rn(alpha, ia));
    }
    alpha = fminf(fmaxf(alpha, 0.0f), 1.0f);
    const uint8_t q[3] = {(uint8_t)__fmul_rn(fminf(fmaxf(cr, 0.0f), 1.0f), 255.0f),
                          (uint8_t)__fmul_rn(fminf(fmaxf(cg, 0.0f), 1.0f), 255.0f),
                          (uint8_t)__fmul_rn(fminf(fmaxf(cb, 0.0f), 1.0f), 255.0f)};
    if (out_alpha) out_alpha[i] = alpha;
    const float ia = __fsub_rn(1.0f, alpha);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      if (out_rgb) out_rgb[3 * i + c] = q[c];
      const float gf = __fdiv_rn((float)q[c], 255.0f);
      if (out_blend) {
        const float bf = __fdiv_rn((float)bg_rgb[3 * i + c], 255.0f);
        const float m = __fadd_rn(__fmul_rn(gf, alpha), __fmul_rn(bf, ia));
        out_blend[3 * i + c] = (uint8_t)__fmul_rn(fminf(fmaxf(m, 0.0f), 1.0f), 255.0f);
      }
      if (out_masked) out_masked[3 * i + c] = (uint8_t)__fmul_rn(__fmul_rn(gf, alpha), 255.0f);
    }
  }
}

int grid_for(long long items) {
  long long blocks = (items + 255) / 256;
  if (blocks > 148LL * 8) blocks = 148LL * 8;
  return (int)(blocks < 1 ? 1 : blocks);
}

}  // namespace

extern "C" int vc_composite_by_depth(const void* bg_rgb, const void* bg_depth, const void* bg_mask, const void* fg_rgb,
                                     const void* fg_depth, const void* fg_mask, void* out_rgb, void* out_depth,
                                     void* out_mask_rgb, long long n_pix, void* stream) {
  VC_CHECK_ARG(bg_depth && fg_depth && fg_mask, "vc_composite_by_depth: null depth/mask operand");
  VC_CHECK_ARG(!out_rgb || (bg_rgb && fg_rgb), "vc_composite_by_depth: out_rgb needs bg_rgb and fg_rgb");
  VC_CHECK_ARG(!out_mask_rgb || bg_mask, "vc_composite_by_depth: out_mask_rgb needs bg_mask");
  VC_CHECK_ARG(n_pix >= 0, "vc_composite_by_depth: negative size");
  if (n_pix == 0) return 0;
  composite_kernel<<<grid_for(n_pix), 256, 0, (cudaStream_t)stream>>>(
      (const uint8_t*)bg_rgb, (const float*)bg_depth, (const uint8_t*)bg_mask, (const uint8_t*)fg_rgb,
      (const float*)fg_depth, (const uint8_t*)fg_mask, (uint8_t*)out_rgb, (float*)out_depth, (uint8_t*)out_mask_rgb, n_pix);
  VC_RETURN_LAUNCH_STATUS("vc_composite_by_depth");
}

extern "C" int vc_depth_to_gray(const void* depth, void* out_rgb, long long n_pix, float min_disp, float denom,
                                int normalise, void* stream) {
  VC_CHECK_ARG(depth && out_rgb && n_pix >= 0, "vc_depth_to_gray: bad operand");
  if (n_pix == 0) return 0;
  depth_gray_kernel<<<grid_for(n_pix), 256, 0, (cudaStream_t)stream>>>((const float*)depth, (uint8_t*)out_rgb, n_pix,
                                                                       min_disp, denom, normalise);
  VC_RETURN_LAUNCH_STATUS("vc_depth_to_gray");
}

extern "C" int vc_gaussian_projection(int n_gauss, const float* params /* [n_gauss, 10] host */, int W, int H,
                                      float threshold, const void* bg_rgb, void* max_scratch /* >= 32 u32, device */,
                                      void* out_rgb, void* out_alpha, void* out_blend, void* out_masked, void* stream) {
  VC_CHECK_ARG(n_gauss >= 0 && n_gauss <= MAX_G, "vc_gaussian_projection: 0..%d Gaussians per frame (got %d)", MAX_G, n_gauss);
  VC_CHECK_ARG(W > 0 && H > 0 && max_scratch && (n_gauss == 0 || params), "vc_gaussian_projection: bad operand");
  VC_CHECK_ARG(!out_blend || bg_rgb, "vc_gaussian_projection: out_blend needs bg_rgb");
  GaussFrame f;
  f.n = n_gauss;
  for (int k = 0; k < n_gauss; ++k) {
    const float* q = params + 10 * k;
    f.mx[k] = q[0]; f.my[k] = q[1];
    f.i00[k] = q[2]; f.i01[k] = q[3]; f.i10[k] = q[4]; f.i11[k] = q[5];
    f.coeff[k] = q[6];
    f.r[k] = q[7]; f.g[k] = q[8]; f.b[k] = q[9];
  }
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(max_scratch, 0, MAX_G * sizeof(unsigned int), s);
  if (e != cudaSuccess) {
    vc_host::set_error("vc_gaussian_projection: memset failed: %s", cudaGetErrorString(e));
    return (int)e;
  }
  const long long n_pix = (long long)W * H;
  if (n_gauss > 0) gauss_max_kernel<<<grid_for(n_pix), 256, 0, s>>>(f, W, H, (unsigned int*)max_scratch);
  gauss_composite_kernel<<<grid_for(n_pix), 256, 0, s>>>(f, W, H, (const unsigned int*)max_scratch, threshold,
                                                         (const uint8_t*)bg_rgb, (uint8_t*)out_rgb, (float*)out_alpha,
                                                         (uint8_t*)out_blend, (uint8_t*)out_masked);
  VC_RETURN_LAUNCH_STATUS("vc_gaussian_projection");
}
