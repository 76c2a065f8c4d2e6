// Elementwise / layout kernels either side of the transformer forward inside the denoising loop (SURVEY.md §8f rows
// N1 and N3).  All HBM-bound byte movement: coalesced, vectorised where the layout allows, no shared memory.
//
//   control_latents_kernel : builds the GeoAdapter input [C_lat + 64, F, H', W'] from the VAE latents of the control
//                            videos and the merged mask (pipeline_wan_versecrafter.py:397-489: torch.cat of the
//                            control-video latents on the channel axis, geoada_encode_masks' 8x8 pixel-unshuffle +
//                            nearest-exact temporal resample, geoada_latent's channel concat) in ONE pass.
//   cfg_euler_step_kernel  : classifier-free-guidance combine + flow-matching Euler update of the latents
//                            (pipeline_wan_versecrafter.py:903-909 with FlowMatchEulerDiscreteScheduler.step).
#include "common.cuh"

namespace {

constexpr int MAX_SRC = 8;      // control videos concatenated on the channel axis (the reference uses 4)
constexpr int MAX_DEPTH = 256;  // latent frames (21 at 81 video frames)

struct ControlParams {
  const __nv_bfloat16* lat[MAX_SRC];  // each [c_lat, Fl, Hl, Wl]
  int n_src, c_lat;
  const __nv_bfloat16* mask;          // [T, Hl*8, Wl*8] (channel 0 of the mask video)
  int T, Fl, Hl, Wl;
  int n_ref;                          // leading reference-image frames: latents carry them, mask rows are zero
  int src_frame[MAX_DEPTH];           // nearest-exact source frame of every resampled mask frame
  __nv_bfloat16* out;                 // [n_src*c_lat + 64, Fl, Hl, Wl]
};

// one thread = one output position (f, h, w): copies nothing for the latent part (done by the vector loop below) and
// gathers the 8x8 mask patch -> 64 channel-strided stores (adjacent threads = adjacent w: coalesced both sides)
__global__ void __launch_bounds__(256) control_latents_kernel(const __grid_constant__ ControlParams p) {
  const long long plane = (long long)p.Hl * p.Wl;
  const long long vol = (long long)p.Fl * plane;
  const int c_all = p.n_src * p.c_lat;
  // (1) latent channels: straight 16-byte copies, source s -> channels [s*c_lat, (s+1)*c_lat)
  const long long vec_per_src = (long long)p.c_lat * vol / 8;
  const long long n_vec = vec_per_src * p.n_src;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += (long long)gridDim.x * blockDim.x) {
    const int s = (int)(i / vec_per_src);
    const long long j = i - (long long)s * vec_per_src;
    reinterpret_cast<uint4*>(p.out)[i] = __ldg(reinterpret_cast<const uint4*>(p.lat[s]) + j);
  }
  // (2) mask channels 64: out[c_all + i*8 + j, f, h, w] = mask[src_frame[f - n_ref], 8h + i, 8w + j]  (0 for f < n_ref)
  __nv_bfloat16* mout = p.out + (long long)c_all * vol;
  const int Wm = p.Wl * 8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < vol; i += (long long)gridDim.x * blockDim.x) {
    const int w = (int)(i % p.Wl);
    const long long t = i / p.Wl;
    const int h = (int)(t % p.Hl);
    const int f = (int)(t / p.Hl);
    if (f < p.n_ref) {
#pragma unroll 8
      for (int c = 0; c < 64; ++c) mout[(long long)c * vol + i] = __float2bfloat16(0.0f);
      continue;
    }
    const __nv_bfloat16* src = p.mask + ((long long)p.src_frame[f - p.n_ref] * p.Hl * 8 + (long long)h * 8) * Wm + (long long)w * 8;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(src + (long long)r * Wm));
      const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        __nv_bfloat162 pr = *reinterpret_cast<const __nv_bfloat162*>(&u[q]);
        mout[(long long)(r * 8 + 2 * q) * vol + i] = pr.x;
        mout[(long long)(r * 8 + 2 * q + 1) * vol + i] = pr.y;
      }
    }
  }
}

// latents_out = bf16( fp32(latents) + bf16(dt * guided) ),  guided = bf16(u + bf16(g * bf16(c - u)))
// (scheduler.step: sample.float() + dt * model_output with a 0-dim fp32 dt and a bf16 model_output -> the product is
// a bf16 tensor, the sum fp32, the result cast back to model_output.dtype)
// u / c = unconditional / text-conditional halves of the CFG batch (pipeline:905-906); each bf16() is a rounding the
// reference's op-by-op bf16 tensor arithmetic performs.  guidance == 1 (no CFG): guided = the single prediction.
__global__ void __launch_bounds__(256) cfg_euler_step_kernel(const __nv_bfloat16* __restrict__ pred_uncond,
                                                            const __nv_bfloat16* __restrict__ pred_text,
                                                            const __nv_bfloat16* __restrict__ latents,
                                                            __nv_bfloat16* __restrict__ latents_out,
                                                            __nv_bfloat16* __restrict__ guided_out, long long n_vec,
                                                            float guidance, float dt, int use_cfg) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += (long long)gridDim.x * blockDim.x) {
    const uint4 uu = __ldg(reinterpret_cast<const uint4*>(pred_uncond) + i);
    const uint4 xx = __ldg(reinterpret_cast<const uint4*>(latents) + i);
    uint4 cc = uu;
    if (use_cfg) cc = __ldg(reinterpret_cast<const uint4*>(pred_text) + i);
    const uint32_t ua[4] = {uu.x, uu.y, uu.z, uu.w}, ca[4] = {cc.x, cc.y, cc.z, cc.w}, xa[4] = {xx.x, xx.y, xx.z, xx.w};
    uint32_t oa[4], ga[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 u = vc::unpack_bf16x2(ua[q]), c = vc::unpack_bf16x2(ca[q]), x = vc::unpack_bf16x2(xa[q]);
      float gx = u.x, gy = u.y;
      if (use_cfg) {
        gx = vc::bf16_round(u.x + vc::bf16_round(guidance * vc::bf16_round(c.x - u.x)));
        gy = vc::bf16_round(u.y + vc::bf16_round(guidance * vc::bf16_round(c.y - u.y)));
      }
      ga[q] = vc::pack_bf16x2(gx, gy);
      oa[q] = vc::pack_bf16x2(__fadd_rn(x.x, vc::bf16_round(__fmul_rn(dt, gx))),
                              __fadd_rn(x.y, vc::bf16_round(__fmul_rn(dt, gy))));
    }
    reinterpret_cast<uint4*>(latents_out)[i] = make_uint4(oa[0], oa[1], oa[2], oa[3]);
    if (guided_out) reinterpret_cast<uint4*>(guided_out)[i] = make_uint4(ga[0], ga[1], ga[2], ga[3]);
  }
}

// CFG combine + flow-matching UniPC (predict-x0, bh2, order <= 2) corrector and predictor in one pass over the latents.
// Per element, fp32 (state m0/m1/last are fp32; every product/sum is a separate rounding, no contraction, so a torch
// fp32 restatement with the same op order agrees to the last bit given the same scalars):
//   v   = CFG(u, c)                                   (bf16 rounding points of the pipeline's expression)
//   x0  = x - sigma * v                               (flow prediction -> data prediction)
//   xc  = corr ? (c_x*last - c_m0*m0) - c_bh*(c_rho0*((m1-m0)/c_rk) + c_rhot*(x0-m0)) : x
//   out = (p_x*xc - p_m0*x0) - p_bh*(0.5*((m0-x0)/p_rk))        [last term only at predictor order 2]
// The x0 history rotates on the host: x0_out becomes m0, the old m0 becomes m1; last_out = xc.
struct UniPCCoef {
  float sigma;
  int corr_order;  // 0 = no corrector (first step)
  float c_x, c_m0, c_bh, c_rho0, c_rk, c_rhot;
  int pred_order;
  float p_x, p_m0, p_bh, p_rk;
};

__global__ void __launch_bounds__(256) cfg_unipc_step_kernel(const __nv_bfloat16* __restrict__ pred_uncond,
                                                            const __nv_bfloat16* __restrict__ pred_text,
                                                            const __nv_bfloat16* __restrict__ latents,
                                                            const float* __restrict__ last, const float* __restrict__ m0,
                                                            const float* __restrict__ m1,
                                                            __nv_bfloat16* __restrict__ latents_out,
                                                            float* __restrict__ x0_out, float* __restrict__ last_out,
                                                            long long n_vec, float guidance, int use_cfg, UniPCCoef k) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += (long long)gridDim.x * blockDim.x) {
    const uint2 uu = __ldg(reinterpret_cast<const uint2*>(pred_uncond) + i);
    const uint2 xx = __ldg(reinterpret_cast<const uint2*>(latents) + i);
    uint2 cc = uu;
    if (use_cfg) cc = __ldg(reinterpret_cast<const uint2*>(pred_text) + i);
    const float2 u01 = vc::unpack_bf16x2(uu.x), u23 = vc::unpack_bf16x2(uu.y);
    const float2 c01 = vc::unpack_bf16x2(cc.x), c23 = vc::unpack_bf16x2(cc.y);
    const float2 x01 = vc::unpack_bf16x2(xx.x), x23 = vc::unpack_bf16x2(xx.y);
    const float u[4] = {u01.x, u01.y, u23.x, u23.y}, c[4] = {c01.x, c01.y, c23.x, c23.y};
    const float x[4] = {x01.x, x01.y, x23.x, x23.y};
    float4 lv = make_float4(0.f, 0.f, 0.f, 0.f), m0v = lv, m1v = lv;
    if (k.corr_order >= 1) {
      lv = __ldg(reinterpret_cast<const float4*>(last) + i);
      m0v = __ldg(reinterpret_cast<const float4*>(m0) + i);
    } else if (k.pred_order == 2) {
      m0v = __ldg(reinterpret_cast<const float4*>(m0) + i);
    }
    if (k.corr_order == 2) m1v = __ldg(reinterpret_cast<const float4*>(m1) + i);
    const float la[4] = {lv.x, lv.y, lv.z, lv.w}, m0a[4] = {m0v.x, m0v.y, m0v.z, m0v.w};
    const float m1a[4] = {m1v.x, m1v.y, m1v.z, m1v.w};
    float x0a[4], xca[4], oa[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float v = u[q];
      if (use_cfg) v = vc::bf16_round(u[q] + vc::bf16_round(guidance * vc::bf16_round(c[q] - u[q])));
      const float x0 = __fsub_rn(x[q], __fmul_rn(k.sigma, v));
      float xc = x[q];
      if (k.corr_order >= 1) {
        const float xt = __fsub_rn(__fmul_rn(k.c_x, la[q]), __fmul_rn(k.c_m0, m0a[q]));
        float res = __fmul_rn(k.c_rhot, __fsub_rn(x0, m0a[q]));
        if (k.corr_order == 2) res = __fadd_rn(__fmul_rn(k.c_rho0, __fdiv_rn(__fsub_rn(m1a[q], m0a[q]), k.c_rk)), res);
        xc = __fsub_rn(xt, __fmul_rn(k.c_bh, res));
      }
      float o = __fsub_rn(__fmul_rn(k.p_x, xc), __fmul_rn(k.p_m0, x0));
      if (k.pred_order == 2)
        o = __fsub_rn(o, __fmul_rn(k.p_bh, __fmul_rn(0.5f, __fdiv_rn(__fsub_rn(m0a[q], x0), k.p_rk))));
      x0a[q] = x0; xca[q] = xc; oa[q] = o;
    }
    reinterpret_cast<float4*>(x0_out)[i] = make_float4(x0a[0], x0a[1], x0a[2], x0a[3]);
    reinterpret_cast<float4*>(last_out)[i] = make_float4(xca[0], xca[1], xca[2], xca[3]);
    reinterpret_cast<uint2*>(latents_out)[i] = make_uint2(vc::pack_bf16x2(oa[0], oa[1]), vc::pack_bf16x2(oa[2], oa[3]));
  }
}

int grid_for(long long work_items, int threads) {
  long long blocks = (work_items + threads - 1) / threads;
  const long long cap = 148LL * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

}  // namespace

extern "C" int vc_control_latents(const void* const* latents, int n_src, int c_lat, const void* mask, int T, int Fl,
                                  int Hl, int Wl, int n_ref, const int* src_frame, void* out, void* stream) {
  VC_CHECK_ARG(latents && mask && out && src_frame, "vc_control_latents: null operand");
  VC_CHECK_ARG(n_src >= 1 && n_src <= MAX_SRC && c_lat >= 1, "vc_control_latents: 1..%d latent sources (got %d)", MAX_SRC,
               n_src);
  VC_CHECK_ARG(T >= 1 && Fl >= 1 && Hl >= 1 && Wl >= 1 && n_ref >= 0 && n_ref < Fl && Fl - n_ref <= MAX_DEPTH,
               "vc_control_latents: bad sizes T=%d Fl=%d Hl=%d Wl=%d n_ref=%d", T, Fl, Hl, Wl, n_ref);
  const long long vol = (long long)Fl * Hl * Wl;
  VC_CHECK_ARG(((long long)c_lat * vol) % 8 == 0, "vc_control_latents: c_lat*F*H*W must be a multiple of 8");
  ControlParams p;
  uintptr_t align = reinterpret_cast<uintptr_t>(mask) | reinterpret_cast<uintptr_t>(out);
  for (int s = 0; s < n_src; ++s) {
    VC_CHECK_ARG(latents[s], "vc_control_latents: null latent source %d", s);
    p.lat[s] = (const __nv_bfloat16*)latents[s];
    align |= reinterpret_cast<uintptr_t>(latents[s]);
  }
  VC_CHECK_ARG(align % 16 == 0, "vc_control_latents: operands must be 16-byte aligned");
  for (int f = 0; f < Fl - n_ref; ++f) {
    VC_CHECK_ARG(src_frame[f] >= 0 && src_frame[f] < T, "vc_control_latents: src_frame[%d] = %d outside [0, %d)", f,
                 src_frame[f], T);
    p.src_frame[f] = src_frame[f];
  }
  p.n_src = n_src; p.c_lat = c_lat;
  p.mask = (const __nv_bfloat16*)mask;
  p.T = T; p.Fl = Fl; p.Hl = Hl; p.Wl = Wl; p.n_ref = n_ref;
  p.out = (__nv_bfloat16*)out;
  control_latents_kernel<<<grid_for(vol, 256), 256, 0, (cudaStream_t)stream>>>(p);
  VC_RETURN_LAUNCH_STATUS("vc_control_latents");
}

extern "C" int vc_cfg_euler_step(const void* pred_uncond, const void* pred_text, const void* latents, void* latents_out,
                                 void* guided_out, long long n, float guidance, float dt, void* stream) {
  VC_CHECK_ARG(pred_uncond && latents && latents_out, "vc_cfg_euler_step: null operand");
  VC_CHECK_ARG(n >= 0 && n % 8 == 0, "vc_cfg_euler_step: element count must be a multiple of 8 (got %lld)", n);
  VC_CHECK_ARG((reinterpret_cast<uintptr_t>(pred_uncond) | reinterpret_cast<uintptr_t>(pred_text) |
                reinterpret_cast<uintptr_t>(latents) | reinterpret_cast<uintptr_t>(latents_out) |
                reinterpret_cast<uintptr_t>(guided_out)) % 16 == 0,
               "vc_cfg_euler_step: operands must be 16-byte aligned");
  if (n == 0) return 0;
  cfg_euler_step_kernel<<<grid_for(n / 8, 256), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)pred_uncond, (const __nv_bfloat16*)pred_text, (const __nv_bfloat16*)latents,
      (__nv_bfloat16*)latents_out, (__nv_bfloat16*)guided_out, n / 8, guidance, dt, pred_text != nullptr ? 1 : 0);
  VC_RETURN_LAUNCH_STATUS("vc_cfg_euler_step");
}

extern "C" int vc_cfg_unipc_step(const void* pred_uncond, const void* pred_text, const void* latents, const void* last,
                                 const void* m0, const void* m1, void* latents_out, void* x0_out, void* last_out,
                                 long long n, float guidance, const float* coef, void* stream) {
  VC_CHECK_ARG(pred_uncond && latents && latents_out && x0_out && last_out && coef, "vc_cfg_unipc_step: null operand");
  VC_CHECK_ARG(n >= 0 && n % 4 == 0, "vc_cfg_unipc_step: element count must be a multiple of 4 (got %lld)", n);
  UniPCCoef k;
  k.sigma = coef[0];
  k.corr_order = (int)coef[1];
  k.c_x = coef[2]; k.c_m0 = coef[3]; k.c_bh = coef[4]; k.c_rho0 = coef[5]; k.c_rk = coef[6]; k.c_rhot = coef[7];
  k.pred_order = (int)coef[8];
  k.p_x = coef[9]; k.p_m0 = coef[10]; k.p_bh = coef[11]; k.p_rk = coef[12];
  VC_CHECK_ARG(k.corr_order >= 0 && k.corr_order <= 2 && k.pred_order >= 1 && k.pred_order <= 2,
               "vc_cfg_unipc_step: corrector order 0..2 and predictor order 1..2 (got %d, %d)", k.corr_order, k.pred_order);
  VC_CHECK_ARG(k.corr_order == 0 || (last && m0), "vc_cfg_unipc_step: the corrector needs `last` and `m0`");
  VC_CHECK_ARG(k.corr_order < 2 || m1, "vc_cfg_unipc_step: the order-2 corrector needs `m1`");
  VC_CHECK_ARG(k.pred_order < 2 || k.corr_order >= 1 || m0, "vc_cfg_unipc_step: the order-2 predictor needs `m0`");
  VC_CHECK_ARG((reinterpret_cast<uintptr_t>(pred_uncond) | reinterpret_cast<uintptr_t>(pred_text) |
                reinterpret_cast<uintptr_t>(latents) | reinterpret_cast<uintptr_t>(latents_out)) % 8 == 0 &&
               (reinterpret_cast<uintptr_t>(last) | reinterpret_cast<uintptr_t>(m0) | reinterpret_cast<uintptr_t>(m1) |
                reinterpret_cast<uintptr_t>(x0_out) | reinterpret_cast<uintptr_t>(last_out)) % 16 == 0,
               "vc_cfg_unipc_step: bf16 operands must be 8-byte and fp32 operands 16-byte aligned");
  if (n == 0) return 0;
  cfg_unipc_step_kernel<<<grid_for(n / 4, 256), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)pred_uncond, (const __nv_bfloat16*)pred_text, (const __nv_bfloat16*)latents,
      (const float*)last, (const float*)m0, (const float*)m1, (__nv_bfloat16*)latents_out, (float*)x0_out,
      (float*)last_out, n / 4, guidance, pred_text != nullptr ? 1 : 0, k);
  VC_RETURN_LAUNCH_STATUS("vc_cfg_unipc_step");
}
