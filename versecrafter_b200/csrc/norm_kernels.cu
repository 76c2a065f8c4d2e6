// HBM-bound row kernels of the DiT block:
//   vc_ln_modulate     : out = LN(x) * (1 + scale[b]) + shift[b]   or   LN(x) * gamma + beta
//                        reference: WanLayerNorm + modulate, wan_transformer3d.py:326-346, :591-592, :603-604, :600
//   vc_rmsnorm_rope_qk : in place on the fused QKV rows: q,k <- RoPE(RMSNorm_D(q|k) * g)
//                        reference: WanRMSNorm :307-323 (over the FULL hidden dim), rope_apply :143-172
// One warp owns one row; the row lives in registers between the statistics pass and the write (one HBM read and
// one HBM write per element), 16-byte vector accesses, warp-shuffle reductions, no shared memory.
// bf16 rounding points follow the reference's op-by-op autocast arithmetic.
#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace {

using vc::bf16_round;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ uint4 ld_stream(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];\n"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
// same, for buffers this kernel also writes (in-place RMSNorm/RoPE): no .nc
__device__ __forceinline__ uint4 ld_stream_rw(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];\n"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}
__device__ __forceinline__ void st_stream(uint4* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1, %2, %3, %4};\n" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}

__device__ __forceinline__ void unpack8(const uint4& u, float* f) {
  float2 a = vc::unpack_bf16x2(u.x), b = vc::unpack_bf16x2(u.y), c = vc::unpack_bf16x2(u.z), d = vc::unpack_bf16x2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 u;
  u.x = vc::pack_bf16x2(f[0], f[1]); u.y = vc::pack_bf16x2(f[2], f[3]);
  u.z = vc::pack_bf16x2(f[4], f[5]); u.w = vc::pack_bf16x2(f[6], f[7]);
  return u;
}

// ----------------------------------------------------------------------------------------------------
// LayerNorm (+ modulate | affine).  VPL = 16-byte vectors per lane (row length D <= VPL*32*8).
// mode 0: out = bf(bf(bf(ln) * bf(1 + scale)) + shift), shift/scale bf16 [batch, D] (row stride mod_stride)
// mode 1: out = bf(ln * gamma + beta)   (fp32 affine inside layer_norm, one rounding; gamma/beta bf16)
// ----------------------------------------------------------------------------------------------------
template <int VPL>
__global__ void __launch_bounds__(256) ln_modulate_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out,
                                                         const __nv_bfloat16* __restrict__ p0,  // shift | beta
                                                         const __nv_bfloat16* __restrict__ p1,  // scale | gamma
                                                         long long mod_stride, long long rows, long long rows_per_batch,
                                                         int D, long long ldx, long long ldo, float eps, int mode) {
  const int lane = threadIdx.x & 31;
  const int nvec = D >> 3;
  const long long warps_total = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows; row += warps_total) {
    const uint4* xr = reinterpret_cast<const uint4*>(x + row * ldx);
    uint4 raw[VPL];
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      int v = lane + 32 * i;
      raw[i] = (v < nvec) ? ld_stream(xr + v) : make_uint4(0, 0, 0, 0);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      float f[8];
      unpack8(raw[i], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) s += f[j];
    }
    const float mean = warp_sum(s) / (float)D;
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      if (lane + 32 * i < nvec) {
        float f[8];
        unpack8(raw[i], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float d = f[j] - mean;
          ss += d * d;
        }
      }
    }
    const float rstd = rsqrtf(warp_sum(ss) / (float)D + eps);
    const long long b = row / rows_per_batch;
    const uint4* q0 = reinterpret_cast<const uint4*>(p0 + (mode == 0 ? b * mod_stride : 0));
    const uint4* q1 = reinterpret_cast<const uint4*>(p1 + (mode == 0 ? b * mod_stride : 0));
    uint4* orow = reinterpret_cast<uint4*>(out + row * ldo);
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      int v = lane + 32 * i;
      if (v < nvec) {
        float f[8], a[8], c[8];
        unpack8(raw[i], f);
        unpack8(__ldg(q0 + v), a);
        unpack8(__ldg(q1 + v), c);
        if (mode == 0) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float ln = bf16_round((f[j] - mean) * rstd);
            float sc = bf16_round(1.0f + c[j]);
            f[j] = bf16_round(ln * sc) + a[j];
          }
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = (f[j] - mean) * rstd * c[j] + a[j];
        }
        st_stream(orow + v, pack8(f));
      }
    }
  }
}

// ----------------------------------------------------------------------------------------------------
// QK RMSNorm (full hidden dim) + 3-axis RoPE, in place.  Each warp handles the q row then the k row of a token.
// qkv row layout: [q(D) | k(D) | v(D)] with row stride ld (elements).  Pairs are adjacent elements (2j, 2j+1).
// rope table: float2 (cos, sin) rows [F*nf | H*nh | W*nw] flattened: tab_f[f][nf], tab_h[h][nh], tab_w[w][nw].
// token position = pos_offset + (row % rows_per_batch); (f,h,w) from the global position, w fastest.
// Rows whose position >= F*H*W pass through un-rotated (reference :168).
// ----------------------------------------------------------------------------------------------------
template <int VPL>
__global__ void __launch_bounds__(256) rmsnorm_rope_kernel(__nv_bfloat16* __restrict__ qkv, long long ld, long long rows,
                                                          long long rows_per_batch, long long pos_offset, int D,
                                                          const __nv_bfloat16* __restrict__ gq,
                                                          const __nv_bfloat16* __restrict__ gk,
                                                          const float2* __restrict__ tab, int Fg, int Hg, int Wg,
                                                          int nf, int nh, int nw, float eps, int do_rope, int n_mats) {
  const int lane = threadIdx.x & 31;
  const int nvec = D >> 3;
  const int half_d = nf + nh + nw;  // complex pairs per head (64)
  const long long warps_total = (long long)gridDim.x * (blockDim.x >> 5);
  const float2* tab_f = tab;
  const float2* tab_h = tab + (long long)Fg * nf;
  const float2* tab_w = tab_h + (long long)Hg * nh;
  for (long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows; row += warps_total) {
    const long long pos = pos_offset + (row % rows_per_batch);
    const bool rot = do_rope && pos < (long long)Fg * Hg * Wg;
    int pw = 0, ph = 0, pf = 0;
    if (rot) {
      pw = (int)(pos % Wg);
      ph = (int)((pos / Wg) % Hg);
      pf = (int)(pos / ((long long)Wg * Hg));
    }
#pragma unroll 1
    for (int which = 0; which < n_mats; ++which) {
      uint4* xr = reinterpret_cast<uint4*>(qkv + row * ld + (long long)which * D);
      const uint4* gr = reinterpret_cast<const uint4*>(which == 0 ? gq : gk);
      uint4 raw[VPL];
#pragma unroll
      for (int i = 0; i < VPL; ++i) {
        int v = lane + 32 * i;
        raw[i] = (v < nvec) ? *(xr + v) : make_uint4(0, 0, 0, 0);
      }
      float ss = 0.f;
#pragma unroll
      for (int i = 0; i < VPL; ++i) {
        float f[8];
        unpack8(raw[i], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) ss += f[j] * f[j];
      }
      const float r = bf16_round(rsqrtf(warp_sum(ss) / (float)D + eps));
#pragma unroll
      for (int i = 0; i < VPL; ++i) {
        int v = lane + 32 * i;
        if (v < nvec) {
          float f[8], g[8];
          unpack8(raw[i], f);
          unpack8(__ldg(gr + v), g);
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = bf16_round(bf16_round(f[j] * r) * g[j]);
          if (rot) {
            const int j0 = ((v * 8) % (2 * half_d)) >> 1;  // first complex pair of this vector within its head
#pragma unroll
            for (int p = 0; p < 4; ++p) {
              int j = j0 + p;
              float2 cs;
              if (j < nf) cs = __ldg(tab_f + pf * nf + j);
              else if (j < nf + nh) cs = __ldg(tab_h + ph * nh + (j - nf));
              else cs = __ldg(tab_w + pw * nw + (j - nf - nh));
              float a = f[2 * p], b = f[2 * p + 1];
              f[2 * p] = a * cs.x - b * cs.y;
              f[2 * p + 1] = a * cs.y + b * cs.x;
            }
          }
          *(xr + v) = pack8(f);
        }
      }
    }
  }
}

// ----------------------------------------------------------------------------------------------------
// Wide rows (D > 1024): one 128-thread CTA per row, VPT 16-byte vectors per thread (5 at D = 5120), so a thread holds
// 20 data registers instead of 80 and 8+ CTAs stay resident per SM — enough loads in flight to cover HBM latency.
// Cross-warp sums go through two tiny shared arrays (one __syncthreads each).
// ----------------------------------------------------------------------------------------------------
constexpr int ROW_NT = 128;

// Row -> (batch, position) -> (f, h, w) in 32-bit unsigned arithmetic: rows, positions and the grid volume are < 2^31 (checked by
// the launchers), and nvcc's 64-bit division costs ~5x the issue slots of the 32-bit one — with five of them per row these
// integer sequences were ~15 % of an issue-bound row kernel.
__device__ __forceinline__ unsigned row_batch(long long row, long long rows_per_batch, unsigned& in_batch) {
  const unsigned r = (unsigned)row, n = (unsigned)rows_per_batch;
  const unsigned b = r / n;
  in_batch = r - b * n;
  return b;
}
__device__ __forceinline__ void split_pos(unsigned pos, int Hg, int Wg, int& pf, int& ph, int& pw) {
  const unsigned t = pos / (unsigned)Wg;
  pw = (int)(pos - t * (unsigned)Wg);
  const unsigned f = t / (unsigned)Hg;
  ph = (int)(t - f * (unsigned)Hg);
  pf = (int)f;
}

__device__ __forceinline__ float block_sum(float v, float* red) {  // red: ROW_NT/32 floats, reused only after a sync
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int i = 0; i < ROW_NT / 32; ++i) t += red[i];
  return t;
}

// Packed-bf16 helpers.  The reference's op-by-op bf16 arithmetic rounds every product / sum of two bf16 values once;
// a product of two bf16 values is exact in fp32, so bf16_round(a * b) == HMUL2.BF16 and bf16_round(a + b) == HADD2.BF16
// (the _rn intrinsics forbid contraction into an fma, which would drop a rounding).  Working on bf16x2 words keeps the
// row kernels off the quarter-rate F2F conversion pipe: ~8 instructions per element instead of ~32.
__device__ __forceinline__ __nv_bfloat162 as_bf162(uint32_t u) { return *reinterpret_cast<__nv_bfloat162*>(&u); }
__device__ __forceinline__ uint32_t as_u32(__nv_bfloat162 v) { return *reinterpret_cast<uint32_t*>(&v); }
__device__ __forceinline__ float2 up2(uint32_t u) { return vc::unpack_bf16x2(u); }
// Packed fp32x2 arithmetic (sm_100 FADD2 / FMUL2 / FFMA2: one issue slot for two fp32 lanes).  Per-lane results are those of
// the scalar instructions (round-to-nearest, ftz as everywhere in this file), so only the ORDER of a reduction can change bits.
__device__ __forceinline__ float2 add2(float2 a, float2 b) {
  float2 d;
  asm("{\n\t.reg .b64 ra, rb, rd;\n\tmov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\t"
      "add.rn.ftz.f32x2 rd, ra, rb;\n\tmov.b64 {%0, %1}, rd;\n\t}\n"
      : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return d;
}
__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
  float2 d;
  asm("{\n\t.reg .b64 ra, rb, rd;\n\tmov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\t"
      "mul.rn.ftz.f32x2 rd, ra, rb;\n\tmov.b64 {%0, %1}, rd;\n\t}\n"
      : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return d;
}
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
  float2 d;
  asm("{\n\t.reg .b64 ra, rb, rc, rd;\n\tmov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\tmov.b64 rc, {%6, %7};\n\t"
      "fma.rn.ftz.f32x2 rd, ra, rb, rc;\n\tmov.b64 {%0, %1}, rd;\n\t}\n"
      : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
  return d;
}

// sum of squares of one 16-byte vector into two packed accumulators (even / odd pairs): 3 issue slots per bf16x2 word.
// Every RMSNorm kernel of this file (in place: block, ring; Ulysses producers: scatter, bulk scatter) adds in this order, so
// they stay bit-identical to one another.
__device__ __forceinline__ void vec_sumsq0(const uint4& u, float2& a0, float2& a1) {
  const float2 f0 = up2(u.x), f1 = up2(u.y), f2 = up2(u.z), f3 = up2(u.w);
  a0 = fma2(f0, f0, a0);
  a1 = fma2(f1, f1, a1);
  a0 = fma2(f2, f2, a0);
  a1 = fma2(f3, f3, a1);
}
__device__ __forceinline__ float sumsq_total(const float2& a0, const float2& a1) { return (a0.x + a0.y) + (a1.x + a1.y); }
// one 16-byte vector of q or k: bf(bf(x * r) * g), then the adjacent-pair rotation in fp32 (one rounding at the pack)
__device__ __forceinline__ uint4 rms_rope_vec(const uint4& x, __nv_bfloat162 r2, const uint4& g, bool rot, const float2* cs) {
  const uint32_t xw[4] = {x.x, x.y, x.z, x.w}, gw[4] = {g.x, g.y, g.z, g.w};
  uint32_t ow[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const __nv_bfloat162 t = __hmul2_rn(__hmul2_rn(as_bf162(xw[q]), r2), as_bf162(gw[q]));
    if (rot) {
      const float2 f = __bfloat1622float2(t);
      ow[q] = vc::pack_bf16x2(f.x * cs[q].x - f.y * cs[q].y, f.x * cs[q].y + f.y * cs[q].x);
    } else {
      ow[q] = as_u32(t);
    }
  }
  return make_uint4(ow[0], ow[1], ow[2], ow[3]);
}


// LayerNorm of one row held as VPT 16-byte vectors per thread (ROW_NT threads), shared by the register and the ring kernel.
// Round 2: the row is unpacked to fp32 pairs ONCE and all three passes (sum, centred sum of squares, normalise) run on packed
// fp32x2 instructions: ~260 instead of ~430 issue slots per thread and 10 KB row — at the 1.4 GHz the power cap leaves these
// kernels inside a step they are issue-bound, not HBM-bound (ncu: issue-active 62 % at 1.96 GHz).  Statistics are summed in two
// (even / odd pair) packed accumulators.  Rounding points of the output as before (mode 0 / 1, see the kernels below).
template <int VPT>
__device__ __forceinline__ void ln_row(const uint4 (&raw)[VPT], const uint4 (&sc)[VPT], const uint4 (&sh)[VPT], int nvec,
                                       float inv_d, float eps, int mode, float* red_a, float* red_b, uint4* orow) {
  float2 xf[VPT * 4];
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    xf[i * 4 + 0] = up2(raw[i].x); xf[i * 4 + 1] = up2(raw[i].y);
    xf[i * 4 + 2] = up2(raw[i].z); xf[i * 4 + 3] = up2(raw[i].w);
  }
  float2 a0 = make_float2(0.f, 0.f), a1 = make_float2(0.f, 0.f);
#pragma unroll
  for (int k = 0; k < VPT * 4; k += 2) {
    a0 = add2(a0, xf[k]);
    a1 = add2(a1, xf[k + 1]);
  }
  const float mean = block_sum((a0.x + a0.y) + (a1.x + a1.y), red_a) * inv_d;
  const float2 nm = make_float2(-mean, -mean);
  a0 = make_float2(0.f, 0.f);
  a1 = make_float2(0.f, 0.f);
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    if (threadIdx.x + ROW_NT * i < nvec) {   // masked vectors are zeros: they must not add mean^2
      const float2 d0 = add2(xf[i * 4 + 0], nm), d1 = add2(xf[i * 4 + 1], nm);
      const float2 d2 = add2(xf[i * 4 + 2], nm), d3 = add2(xf[i * 4 + 3], nm);
      a0 = fma2(d0, d0, a0); a1 = fma2(d1, d1, a1);
      a0 = fma2(d2, d2, a0); a1 = fma2(d3, d3, a1);
    }
  }
  const float rstd = rsqrtf(block_sum((a0.x + a0.y) + (a1.x + a1.y), red_b) * inv_d + eps);
  const float2 r2 = make_float2(rstd, rstd);
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + ROW_NT * i;
    if (v < nvec) {
      const uint32_t cw[4] = {sc[i].x, sc[i].y, sc[i].z, sc[i].w};
      const uint32_t aw[4] = {sh[i].x, sh[i].y, sh[i].z, sh[i].w};
      uint32_t ow[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float2 l = mul2(add2(xf[i * 4 + q], nm), r2);   // (x - mean) * rstd, one rounding each as in the scalar form
        if (mode == 0) {
          // bf(bf(bf(ln) * bf(1 + scale)) + shift)
          ow[q] = as_u32(__hadd2_rn(__hmul2_rn(__floats2bfloat162_rn(l.x, l.y), as_bf162(cw[q])), as_bf162(aw[q])));
        } else {
          const float2 o = fma2(l, up2(cw[q]), up2(aw[q]));
          ow[q] = vc::pack_bf16x2(o.x, o.y);
        }
      }
      st_stream(orow + v, make_uint4(ow[0], ow[1], ow[2], ow[3]));
    }
  }
}

template <int VPT>
__global__ void __launch_bounds__(ROW_NT, VPT <= 5 ? 5 : 2) ln_modulate_block_kernel(
    const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ p0,
    const __nv_bfloat16* __restrict__ p1, long long mod_stride, long long rows, long long rows_per_batch, int D,
    long long ldx, long long ldo, float eps, int mode) {
  __shared__ float red_a[ROW_NT / 32], red_b[ROW_NT / 32];
  const int nvec = D >> 3;
  const float inv_d = 1.0f / (float)D;
  // A thread owns the same columns of every row, so the per-column operands stay in registers while the batch index
  // does not change: mode 0: sc = bf16(1 + scale), sh = shift;  mode 1: sc = gamma, sh = beta (all packed bf16x2).
  // (A next-row register prefetch was measured and made no difference once the arithmetic was packed.)
  uint4 raw[VPT], sc[VPT], sh[VPT];
  long long cached_batch = -1;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const uint4* xr = reinterpret_cast<const uint4*>(x + row * ldx);
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + ROW_NT * i;
      raw[i] = (v < nvec) ? ld_stream(xr + v) : make_uint4(0, 0, 0, 0);
    }
    unsigned in_batch_unused;
    const long long batch = (mode == 0) ? (long long)row_batch(row, rows_per_batch, in_batch_unused) : 0;
    if (batch != cached_batch) {
      cached_batch = batch;
      const uint4* q0 = reinterpret_cast<const uint4*>(p0 + batch * mod_stride);
      const uint4* q1 = reinterpret_cast<const uint4*>(p1 + batch * mod_stride);
      const __nv_bfloat162 one2 = __floats2bfloat162_rn(1.0f, 1.0f);
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + ROW_NT * i;
        if (v < nvec) {
          sh[i] = __ldg(q0 + v);
          uint4 c = __ldg(q1 + v);
          if (mode == 0) {
            c.x = as_u32(__hadd2_rn(one2, as_bf162(c.x)));
            c.y = as_u32(__hadd2_rn(one2, as_bf162(c.y)));
            c.z = as_u32(__hadd2_rn(one2, as_bf162(c.z)));
            c.w = as_u32(__hadd2_rn(one2, as_bf162(c.w)));
          }
          sc[i] = c;
        }
      }
    }
    ln_row<VPT>(raw, sc, sh, nvec, inv_d, eps, mode, red_a, red_b, reinterpret_cast<uint4*>(out + row * ldo));
  }
}

// Requires (ROW_NT * 8) % head_dim == 0 when rotating: every vector of a thread then sits at the same offset inside
// its head, so the 4 (cos, sin) pairs are fetched once per row and reused for all VPT vectors and for q and k.
template <int VPT>
__global__ void __launch_bounds__(ROW_NT, VPT <= 5 ? 5 : 4) rmsnorm_rope_block_kernel(
    __nv_bfloat16* __restrict__ qkv, long long ld, long long rows, long long rows_per_batch, long long pos_offset, int D,
    const __nv_bfloat16* __restrict__ gq, const __nv_bfloat16* __restrict__ gk, const float2* __restrict__ tab, int Fg,
    int Hg, int Wg, int nf, int nh, int nw, float eps, int do_rope, int n_mats) {
  __shared__ float red_a[ROW_NT / 32], red_b[ROW_NT / 32];
  const int nvec = D >> 3;
  const float inv_d = 1.0f / (float)D;
  const int half_d = nf + nh + nw;
  // A thread's four (cos, sin) pairs sit at fixed columns j0..j0+3 of the head for every row, so which axis table a pair comes
  // from, its column inside that table and the table's row stride are loop invariants: per row only the table ROW (frame,
  // height or width index) changes.
  const float2* tcol[4];
  int taxis[4], tstride[4];
  {
    const int j0 = do_rope ? (((threadIdx.x * 8) % (2 * half_d)) >> 1) : 0;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      const int j = j0 + p;
      if (j < nf) { taxis[p] = 0; tstride[p] = nf; tcol[p] = tab + j; }
      else if (j < nf + nh) { taxis[p] = 1; tstride[p] = nh; tcol[p] = tab + (long long)Fg * nf + (j - nf); }
      else { taxis[p] = 2; tstride[p] = nw; tcol[p] = tab + (long long)Fg * nf + (long long)Hg * nh + (j - nf - nh); }
    }
  }
  // software pipeline over (row, q|k) segments: the next segment's loads are issued before the current one's
  // reduction, so a CTA always has two segments of HBM reads in flight
  uint4 raw[VPT], nxt[VPT];
  if ((long long)blockIdx.x < rows) {
    const uint4* x0 = reinterpret_cast<const uint4*>(qkv + (long long)blockIdx.x * ld);
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + ROW_NT * i;
      raw[i] = (v < nvec) ? ld_stream_rw(x0 + v) : make_uint4(0, 0, 0, 0);
    }
  }
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    unsigned in_batch;
    row_batch(row, rows_per_batch, in_batch);
    const unsigned pos = (unsigned)pos_offset + in_batch;
    const bool rot = do_rope && pos < (unsigned)(Fg * Hg * Wg);
    float2 cs[4];
    if (rot) {
      int pf, ph, pw;
      split_pos(pos, Hg, Wg, pf, ph, pw);
#pragma unroll
      for (int p = 0; p < 4; ++p) cs[p] = __ldg(tcol[p] + (taxis[p] == 0 ? pf : taxis[p] == 1 ? ph : pw) * tstride[p]);
    }
#pragma unroll 1
    for (int which = 0; which < n_mats; ++which) {
      uint4* xr = reinterpret_cast<uint4*>(qkv + row * ld + (long long)which * D);
      const uint4* gr = reinterpret_cast<const uint4*>(which == 0 ? gq : gk);
      {  // prefetch the next segment: k of this row, else q of this CTA's next row
        const bool same_row = which + 1 < n_mats;
        const long long nrow = same_row ? row : row + gridDim.x;
        if (nrow < rows) {
          const uint4* xn = reinterpret_cast<const uint4*>(qkv + nrow * ld + (same_row ? (long long)(which + 1) * D : 0));
#pragma unroll
          for (int i = 0; i < VPT; ++i) {
            const int v = threadIdx.x + ROW_NT * i;
            nxt[i] = (v < nvec) ? ld_stream_rw(xn + v) : make_uint4(0, 0, 0, 0);
          }
        }
      }
      float2 sq0 = make_float2(0.f, 0.f), sq1 = make_float2(0.f, 0.f);
#pragma unroll
      for (int i = 0; i < VPT; ++i) vec_sumsq0(raw[i], sq0, sq1);
      const float ss = sumsq_total(sq0, sq1);
      const float r = rsqrtf(block_sum(ss, which ? red_b : red_a) * inv_d + eps);
      const __nv_bfloat162 r2 = __floats2bfloat162_rn(r, r);
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + ROW_NT * i;
        if (v < nvec) st_stream(xr + v, rms_rope_vec(raw[i], r2, __ldg(gr + v), rot, cs));
      }
#pragma unroll
      for (int i = 0; i < VPT; ++i) raw[i] = nxt[i];
      if (n_mats == 1) __syncthreads();  // red_a is reused by the next row
    }
  }
}

// ----------------------------------------------------------------------------------------------------
// Row kernels with an asynchronous row ring (round 2).  ncu on the register-resident kernels above (profiles/
// ncu_ln_r01j_summary.txt): 28 % warps active, 49 % issue-active, long-scoreboard bound — one 10 KB row per CTA does not
// keep enough HBM reads in flight once the power cap drops the clock.  Here a CTA keeps RING rows of input in flight in
// shared memory: one thread issues 1-D bulk copies (cp.async.bulk, the TMA engine) two rows ahead of the row being
// reduced; the 128 threads pick their 16-byte vectors out of shared memory when the row's mbarrier completes.  Arithmetic
// and rounding points are those of the kernels above (same helpers), so results are bit-identical.
// ----------------------------------------------------------------------------------------------------
constexpr int RING = 3;

template <int VPT>
__global__ void __launch_bounds__(ROW_NT, VPT <= 5 ? 5 : 2) ln_modulate_ring_kernel(
    const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ p0,
    const __nv_bfloat16* __restrict__ p1, long long mod_stride, long long rows, long long rows_per_batch, int D,
    long long ldx, long long ldo, float eps, int mode) {
  extern __shared__ __align__(16) uint8_t ring_smem[];
  __shared__ float red_a[ROW_NT / 32], red_b[ROW_NT / 32];
  __shared__ __align__(8) uint64_t full[RING];
  const int nvec = D >> 3;
  const uint32_t row_bytes = (uint32_t)D * 2;
  const float inv_d = 1.0f / (float)D;
  const long long my_rows = rows > (long long)blockIdx.x ? (rows - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  if (threadIdx.x == 0) {
    for (int i = 0; i < RING; ++i) vc::mbar_init(&full[i], 1);
    vc::fence_barrier_init();
    for (int i = 0; i < RING - 1 && i < my_rows; ++i) {
      vc::mbar_expect_tx(&full[i], row_bytes);
      vc::bulk_load_1d(ring_smem + (size_t)i * row_bytes, x + ((long long)blockIdx.x + (long long)i * gridDim.x) * ldx, row_bytes,
                       &full[i]);
    }
  }
  __syncthreads();
  uint4 raw[VPT], sc[VPT], sh[VPT];
  long long cached_batch = -1;
  for (long long it = 0; it < my_rows; ++it) {
    const long long row = (long long)blockIdx.x + it * gridDim.x;
    const int slot = (int)(it % RING);
    // refill the slot consumed in the previous iteration (every thread copied it to registers before that iteration's
    // first __syncthreads) with the row RING-1 ahead
    if (threadIdx.x == 0 && it + RING - 1 < my_rows) {
      const int ns = (int)((it + RING - 1) % RING);
      vc::mbar_expect_tx(&full[ns], row_bytes);
      vc::bulk_load_1d(ring_smem + (size_t)ns * row_bytes, x + (row + (long long)(RING - 1) * gridDim.x) * ldx, row_bytes,
                       &full[ns]);
    }
    unsigned in_batch_unused;
    const long long batch = (mode == 0) ? (long long)row_batch(row, rows_per_batch, in_batch_unused) : 0;
    if (batch != cached_batch) {
      cached_batch = batch;
      const uint4* q0 = reinterpret_cast<const uint4*>(p0 + batch * mod_stride);
      const uint4* q1 = reinterpret_cast<const uint4*>(p1 + batch * mod_stride);
      const __nv_bfloat162 one2 = __floats2bfloat162_rn(1.0f, 1.0f);
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + ROW_NT * i;
        if (v < nvec) {
          sh[i] = __ldg(q0 + v);
          uint4 c = __ldg(q1 + v);
          if (mode == 0) {
            c.x = as_u32(__hadd2_rn(one2, as_bf162(c.x)));
            c.y = as_u32(__hadd2_rn(one2, as_bf162(c.y)));
            c.z = as_u32(__hadd2_rn(one2, as_bf162(c.z)));
            c.w = as_u32(__hadd2_rn(one2, as_bf162(c.w)));
          }
          sc[i] = c;
        }
      }
    }
    vc::mbar_wait(&full[slot], (uint32_t)((it / RING) & 1));
    const uint4* xr = reinterpret_cast<const uint4*>(ring_smem + (size_t)slot * row_bytes);
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + ROW_NT * i;
      raw[i] = (v < nvec) ? xr[v] : make_uint4(0, 0, 0, 0);
    }
    ln_row<VPT>(raw, sc, sh, nvec, inv_d, eps, mode, red_a, red_b, reinterpret_cast<uint4*>(out + row * ldo));
  }
}

// RMSNorm (+ RoPE) in place with the same ring; a ring entry is one (row, q|k) segment of D elements.
template <int VPT>
__global__ void __launch_bounds__(ROW_NT, VPT <= 5 ? 5 : 4) rmsnorm_rope_ring_kernel(
    __nv_bfloat16* __restrict__ qkv, long long ld, long long rows, long long rows_per_batch, long long pos_offset, int D,
    const __nv_bfloat16* __restrict__ gq, const __nv_bfloat16* __restrict__ gk, const float2* __restrict__ tab, int Fg,
    int Hg, int Wg, int nf, int nh, int nw, float eps, int do_rope, int n_mats) {
  extern __shared__ __align__(16) uint8_t ring_smem[];
  __shared__ float red_a[ROW_NT / 32], red_b[ROW_NT / 32];
  __shared__ __align__(8) uint64_t full[RING];
  const int nvec = D >> 3;
  const uint32_t row_bytes = (uint32_t)D * 2;
  const float inv_d = 1.0f / (float)D;
  const int half_d = nf + nh + nw;
  const float2* tab_f = tab;
  const float2* tab_h = tab + (long long)Fg * nf;
  const float2* tab_w = tab_h + (long long)Hg * nh;
  const int j0 = do_rope ? (((threadIdx.x * 8) % (2 * half_d)) >> 1) : 0;
  const long long my_rows = rows > (long long)blockIdx.x ? (rows - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const long long my_segs = my_rows * n_mats;
  // segment g of this CTA -> global address
  auto seg_ptr = [&](long long g) {
    const long long r = (long long)blockIdx.x + (g / n_mats) * gridDim.x;
    return qkv + r * ld + (g % n_mats) * (long long)D;
  };
  if (threadIdx.x == 0) {
    for (int i = 0; i < RING; ++i) vc::mbar_init(&full[i], 1);
    vc::fence_barrier_init();
    for (int i = 0; i < RING - 1 && i < my_segs; ++i) {
      vc::mbar_expect_tx(&full[i], row_bytes);
      vc::bulk_load_1d(ring_smem + (size_t)i * row_bytes, seg_ptr(i), row_bytes, &full[i]);
    }
  }
  __syncthreads();
  uint4 raw[VPT];
  float2 cs[4];
  bool rot = false;
  for (long long g = 0; g < my_segs; ++g) {
    const int which = (int)(g % n_mats);
    const long long row = (long long)blockIdx.x + (g / n_mats) * gridDim.x;
    const int slot = (int)(g % RING);
    if (threadIdx.x == 0 && g + RING - 1 < my_segs) {
      const int ns = (int)((g + RING - 1) % RING);
      vc::mbar_expect_tx(&full[ns], row_bytes);
      vc::bulk_load_1d(ring_smem + (size_t)ns * row_bytes, seg_ptr(g + RING - 1), row_bytes, &full[ns]);
    }
    if (which == 0) {
      unsigned in_batch;
      row_batch(row, rows_per_batch, in_batch);
      const unsigned pos = (unsigned)pos_offset + in_batch;
      rot = do_rope && pos < (unsigned)(Fg * Hg * Wg);
      if (rot) {
        int pf, ph, pw;
        split_pos(pos, Hg, Wg, pf, ph, pw);
#pragma unroll
        for (int p = 0; p < 4; ++p) {
          const int j = j0 + p;
          if (j < nf) cs[p] = __ldg(tab_f + pf * nf + j);
          else if (j < nf + nh) cs[p] = __ldg(tab_h + ph * nh + (j - nf));
          else cs[p] = __ldg(tab_w + pw * nw + (j - nf - nh));
        }
      }
    }
    vc::mbar_wait(&full[slot], (uint32_t)((g / RING) & 1));
    const uint4* xs = reinterpret_cast<const uint4*>(ring_smem + (size_t)slot * row_bytes);
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + ROW_NT * i;
      raw[i] = (v < nvec) ? xs[v] : make_uint4(0, 0, 0, 0);
    }
    float2 sq0 = make_float2(0.f, 0.f), sq1 = make_float2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < VPT; ++i) vec_sumsq0(raw[i], sq0, sq1);
    const float ss = sumsq_total(sq0, sq1);
    const float r = rsqrtf(block_sum(ss, (g & 1) ? red_b : red_a) * inv_d + eps);
    const __nv_bfloat162 r2 = __floats2bfloat162_rn(r, r);
    uint4* xr = reinterpret_cast<uint4*>(qkv + row * ld + (long long)which * D);
    const uint4* gr = reinterpret_cast<const uint4*>(which == 0 ? gq : gk);
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + ROW_NT * i;
      if (v < nvec) st_stream(xr + v, rms_rope_vec(raw[i], r2, __ldg(gr + v), rot, cs));
    }
  }
}

// ----------------------------------------------------------------------------------------------------
// Ulysses producer: RMSNorm + RoPE of q and k (as above) and a pass-through of v, with the
// result STORED DIRECTLY INTO THE RECEIVE BUFFERS OF THE PEER GPUs over NVLink (peer-mapped pointers) instead of
// in place: rank p receives heads [p*Hp, (p+1)*Hp) of every token.  Destination layout at each peer:
//   recv[which(q,k,v)][b][global token][local head][head_dim]      (bf16)
// so the all-to-all "scatter heads / gather sequence" of the reference's usp_attn_forward (xfuser SeqAllToAll4D,
// not in the reference tree) happens inside this kernel's epilogue: no send-layout copy, no NCCL call, no
// receive-layout copy.  A cross-GPU barrier on the stream (host side) orders it against the attention kernel.
// ----------------------------------------------------------------------------------------------------
struct PeerPtrs {
  __nv_bfloat16* p[8];
};

template <int VPT>
__global__ void __launch_bounds__(ROW_NT, VPT <= 5 ? 8 : 7) rmsnorm_rope_scatter_kernel(
    const __nv_bfloat16* __restrict__ qkv, long long ld, long long rows, long long rows_per_batch, long long pos_offset,
    int D, const __nv_bfloat16* __restrict__ gq, const __nv_bfloat16* __restrict__ gk, const float2* __restrict__ tab,
    int Fg, int Hg, int Wg, int nf, int nh, int nw, float eps, PeerPtrs peers, int heads_per_peer, long long l_total,
    int n_batch, int which_mask) {
  __shared__ float red[2][ROW_NT / 32];  // alternate per reduction: a buffer is rewritten two block syncs after its reads
  int n_red = 0;
  const int nvec = D >> 3;
  const float inv_d = 1.0f / (float)D;
  const int half_d = nf + nh + nw;
  const int vec_per_head = half_d >> 2;  // head_dim / 8
  const float2* tab_f = tab;
  const float2* tab_h = tab + (long long)Fg * nf;
  const float2* tab_w = tab_h + (long long)Hg * nh;
  const int j0 = ((threadIdx.x * 8) % (2 * half_d)) >> 1;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    unsigned in_batch;
    const long long b = row_batch(row, rows_per_batch, in_batch);
    const long long pos = (long long)((unsigned)pos_offset + in_batch);
    const bool rot = pos < (long long)Fg * Hg * Wg;
    float2 cs[4];
    if (rot) {
      int pf, ph, pw;
      split_pos((unsigned)pos, Hg, Wg, pf, ph, pw);
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        const int j = j0 + p;
        if (j < nf) cs[p] = __ldg(tab_f + pf * nf + j);
        else if (j < nf + nh) cs[p] = __ldg(tab_h + ph * nh + (j - nf));
        else cs[p] = __ldg(tab_w + pw * nw + (j - nf - nh));
      }
    }
#pragma unroll 1
    for (int which = 0; which < 3; ++which) {
      if (!((which_mask >> which) & 1)) continue;  // the caller scatters q, k, v in separate launches to overlap them
      const uint4* xr = reinterpret_cast<const uint4*>(qkv + row * ld + (long long)which * D);
      const uint4* gr = reinterpret_cast<const uint4*>(which == 0 ? gq : gk);
      // element offset of (which, b, pos, local head 0, 0) inside a peer's receive buffer
      const long long dst_row = (((long long)which * n_batch + b) * l_total + pos) * heads_per_peer;
      uint4 raw[VPT];
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + ROW_NT * i;
        raw[i] = (v < nvec) ? ld_stream(xr + v) : make_uint4(0, 0, 0, 0);
      }
      __nv_bfloat162 r2 = __floats2bfloat162_rn(1.0f, 1.0f);
      if (which < 2) {
        float2 sq0 = make_float2(0.f, 0.f), sq1 = make_float2(0.f, 0.f);
#pragma unroll
        for (int i = 0; i < VPT; ++i) vec_sumsq0(raw[i], sq0, sq1);
        const float ss = sumsq_total(sq0, sq1);
        const float r = rsqrtf(block_sum(ss, red[n_red & 1]) * inv_d + eps);
        ++n_red;
        r2 = __floats2bfloat162_rn(r, r);
      }
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + ROW_NT * i;
        if (v < nvec) {
          uint4 outv = raw[i];
          if (which < 2) outv = rms_rope_vec(raw[i], r2, __ldg(gr + v), rot, cs);
          const int head = v / vec_per_head;
          const int peer = head / heads_per_peer;
          const int hl = head - peer * heads_per_peer;
          __nv_bfloat16* dst = peers.p[peer] + ((dst_row + hl) * vec_per_head + (v - head * vec_per_head)) * 8;
          *reinterpret_cast<uint4*>(dst) = outv;
        }
      }
    }
  }
}

// Same producer with the stores handed to the TMA engine (round 2): results are staged in shared memory and leave as ONE
// bulk store (cp.async.bulk shared -> global) per (row, matrix, peer) — a peer's heads are contiguous both in the row and
// in its receive block, heads_per_peer * 256 bytes — instead of a 16-byte st.global from every thread.  The per-thread
// stores put tens of thousands of NVLink-bound requests into the SM's load/store path; when this kernel runs on a side
// stream UNDER a projection GEMM (sp.ulysses_qkv_attention_overlapped) those requests sit in front of the GEMM epilogue's
// own stores.  Bulk stores queue in the TMA unit and drain at NVLink speed without occupying warps.  Two staging slots of
// D*2 bytes (20 KB at D = 5120) fit beside the GEMM's 193 KB of shared memory.  Arithmetic identical: bit-identical results.
template <int VPT>
__global__ void __launch_bounds__(ROW_NT, 8) rmsnorm_rope_scatter_bulk_kernel(
    const __nv_bfloat16* __restrict__ qkv, long long ld, long long rows, long long rows_per_batch, long long pos_offset,
    int D, const __nv_bfloat16* __restrict__ gq, const __nv_bfloat16* __restrict__ gk, const float2* __restrict__ tab,
    int Fg, int Hg, int Wg, int nf, int nh, int nw, float eps, PeerPtrs peers, int n_peers, int heads_per_peer,
    long long l_total, int n_batch, int which_mask) {
  extern __shared__ __align__(16) uint8_t stage[];  // 2 output staging slots, D*2 bytes each
  __shared__ float red[2][ROW_NT / 32];
  const int nvec = D >> 3;
  const uint32_t row_bytes = (uint32_t)D * 2;
  const uint32_t peer_bytes = row_bytes / (uint32_t)n_peers;  // heads_per_peer * head_dim * 2
  const float inv_d = 1.0f / (float)D;
  const int half_d = nf + nh + nw;
  const float2* tab_f = tab;
  const float2* tab_h = tab + (long long)Fg * nf;
  const float2* tab_w = tab_h + (long long)Hg * nh;
  const int j0 = ((threadIdx.x * 8) % (2 * half_d)) >> 1;
  int seg = 0;  // segments (row, matrix) handled by this CTA so far: staging slot = seg & 1
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    unsigned in_batch;
    const long long b = row_batch(row, rows_per_batch, in_batch);
    const long long pos = (long long)((unsigned)pos_offset + in_batch);
    const bool rot = pos < (long long)Fg * Hg * Wg;
    float2 cs[4];
    if (rot) {
      int pf, ph, pw;
      split_pos((unsigned)pos, Hg, Wg, pf, ph, pw);
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        const int j = j0 + p;
        if (j < nf) cs[p] = __ldg(tab_f + pf * nf + j);
        else if (j < nf + nh) cs[p] = __ldg(tab_h + ph * nh + (j - nf));
        else cs[p] = __ldg(tab_w + pw * nw + (j - nf - nh));
      }
    }
#pragma unroll 1
    for (int which = 0; which < 3; ++which) {
      if (!((which_mask >> which) & 1)) continue;
      const uint4* xr = reinterpret_cast<const uint4*>(qkv + row * ld + (long long)which * D);
      const uint4* gr = reinterpret_cast<const uint4*>(which == 0 ? gq : gk);
      uint4 raw[VPT];
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + ROW_NT * i;
        raw[i] = (v < nvec) ? ld_stream(xr + v) : make_uint4(0, 0, 0, 0);
      }
      float2 sq0 = make_float2(0.f, 0.f), sq1 = make_float2(0.f, 0.f);
      if (which < 2) {
#pragma unroll
        for (int i = 0; i < VPT; ++i) vec_sumsq0(raw[i], sq0, sq1);
      }
      const float ss = sumsq_total(sq0, sq1);
      // staging slot seg & 1 was last used by segment seg - 2: its bulk stores must have read it out.  Thread 0 waits, the
      // block-wide sync inside block_sum publishes that (for v the sum itself is unused).
      if (threadIdx.x == 0) vc::bulk_wait_group_read<1>();
      const float r = rsqrtf(block_sum(ss, red[seg & 1]) * inv_d + eps);
      const __nv_bfloat162 r2 = __floats2bfloat162_rn(r, r);
      uint4* os = reinterpret_cast<uint4*>(stage + (size_t)(seg & 1) * row_bytes);
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + ROW_NT * i;
        if (v < nvec) os[v] = which < 2 ? rms_rope_vec(raw[i], r2, __ldg(gr + v), rot, cs) : raw[i];
      }
      vc::fence_proxy_async_smem();  // generic-proxy writes of the staging slot -> visible to the TMA engine
      __syncthreads();
      if (threadIdx.x == 0) {
        // element offset of (which, b, pos, local head 0, 0) in a peer's receive buffer; peer p gets row bytes [p*peer_bytes, ..)
        const long long dst = ((((long long)which * n_batch + b) * l_total + pos) * heads_per_peer) * (2LL * half_d);
        for (int p = 0; p < n_peers; ++p)
          vc::bulk_store_1d(peers.p[p] + dst, stage + (size_t)(seg & 1) * row_bytes + (size_t)p * peer_bytes, peer_bytes);
        vc::bulk_commit_group();
      }
      ++seg;
    }
  }
  if (threadIdx.x == 0) vc::bulk_wait_group<0>();
}

// ----------------------------------------------------------------------------------------------------
// patchify: [C, F, H, W] bf16 (one sample) -> A[L, C*4] bf16, K ordered (c, ph, pw)  (SURVEY App. A.1).
// Feeds the tcgen05 GEMM that implements the Conv3d patch embedding (stride == kernel, so it is a pure GEMM).
// thread -> (token, c): reads two 4-byte pairs (rows 2h, 2h+1), writes 8 contiguous bytes.
// ----------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) patchify_kernel(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out,
                                                      int C, int Fr, int H, int W, long long ld_out) {
  const int Hp = H >> 1, Wp = W >> 1;
  const long long L = (long long)Fr * Hp * Wp;
  const long long total = L * C;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    // consecutive threads walk w' first (coalesced 4-byte reads along W), then c, then (f,h)
    const int w = (int)(idx % Wp);
    const long long t1 = idx / Wp;
    const int c = (int)(t1 % C);
    const long long fh = t1 / C;
    const int h = (int)(fh % Hp);
    const int f = (int)(fh / Hp);
    const __nv_bfloat16* src = in + (((long long)c * Fr + f) * H + 2 * h) * W + 2 * w;
    uint32_t r0 = *reinterpret_cast<const uint32_t*>(src);
    uint32_t r1 = *reinterpret_cast<const uint32_t*>(src + W);
    const long long tok = ((long long)f * Hp + h) * Wp + w;
    *reinterpret_cast<uint2*>(out + tok * ld_out + (long long)c * 4) = make_uint2(r0, r1);
  }
}

// unpatchify: y [L, 4*Co] bf16|f32-as-bf16 -> out [Co, F, 2H', 2W'] (wan_transformer3d.py:1127-1150)
__global__ void __launch_bounds__(256) unpatchify_kernel(const __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ out,
                                                        int Co, int Fg, int Hg, int Wg, long long ldy) {
  const long long total = (long long)Co * Fg * Hg * 2 * Wg * 2;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int X = (int)(idx % (2 * Wg));
    long long t = idx / (2 * Wg);
    const int Y = (int)(t % (2 * Hg));
    t /= (2 * Hg);
    const int f = (int)(t % Fg);
    const int c = (int)(t / Fg);
    const long long tok = ((long long)f * Hg + (Y >> 1)) * Wg + (X >> 1);
    out[idx] = y[tok * ldy + ((Y & 1) * 2 + (X & 1)) * Co + c];
  }
}

template <typename F>
int dispatch_vpl(int D, F&& f) {
  const int need = (D / 8 + 31) / 32;
  if (need <= 1) return f(std::integral_constant<int, 1>{});
  if (need <= 2) return f(std::integral_constant<int, 2>{});
  if (need <= 4) return f(std::integral_constant<int, 4>{});
  if (need <= 6) return f(std::integral_constant<int, 6>{});
  if (need <= 8) return f(std::integral_constant<int, 8>{});
  if (need <= 12) return f(std::integral_constant<int, 12>{});
  if (need <= 20) return f(std::integral_constant<int, 20>{});
  if (need <= 32) return f(std::integral_constant<int, 32>{});
  return -1;
}

template <typename F>
int dispatch_vpt(int D, F&& f) {  // block-per-row kernels: vectors per thread at ROW_NT threads
  const int need = (D / 8 + ROW_NT - 1) / ROW_NT;
  if (need <= 2) return f(std::integral_constant<int, 2>{});
  if (need <= 3) return f(std::integral_constant<int, 3>{});
  if (need <= 4) return f(std::integral_constant<int, 4>{});
  if (need <= 5) return f(std::integral_constant<int, 5>{});
  if (need <= 6) return f(std::integral_constant<int, 6>{});
  if (need <= 8) return f(std::integral_constant<int, 8>{});
  return -1;
}

int grid_for_block_rows(long long rows) {
  const long long cap = 148LL * 16;  // up to 16 resident 128-thread CTAs per SM
  return (int)(rows < cap ? (rows < 1 ? 1 : rows) : cap);
}

int grid_for_ring_rows(long long rows, int ctas_per_sm) {  // persistent: every resident CTA walks rows blockIdx.x + i*grid
  const long long cap = (long long)vc_host::sm_count() * ctas_per_sm;
  return (int)(rows < cap ? (rows < 1 ? 1 : rows) : cap);
}

int grid_for_rows(long long rows, int warps_per_cta) {
  long long ctas = (rows + warps_per_cta - 1) / warps_per_cta;
  const long long cap = 148LL * 8;  // 8 resident 256-thread CTAs per SM
  return (int)(ctas < cap ? (ctas < 1 ? 1 : ctas) : cap);
}

}  // namespace

extern "C" {

int vc_ln_modulate(const void* x, long long ldx, void* out, long long ldo, const void* shift_or_beta,
                   const void* scale_or_gamma, long long mod_batch_stride, long long rows, long long rows_per_batch,
                   int D, float eps, int mode, void* stream) {
  VC_CHECK_ARG(x && out && shift_or_beta && scale_or_gamma, "vc_ln_modulate: null pointer");
  VC_CHECK_ARG(D > 0 && D % 8 == 0 && D <= 8192, "vc_ln_modulate: D=%d must be a multiple of 8 and <= 8192", D);
  VC_CHECK_ARG(rows >= 0 && rows_per_batch > 0 && ldx % 8 == 0 && ldo % 8 == 0, "vc_ln_modulate: bad rows/strides");
  VC_CHECK_ARG(rows < (1LL << 31) && rows_per_batch < (1LL << 31), "vc_ln_modulate: rows must be < 2^31");
  const bool force_regs = (mode & 16) != 0;  // schedule bits (A/B, tests): 16 = register-resident kernel, 32 = row ring
  const bool force_ring = (mode & 32) != 0;
  mode &= ~48;
  VC_CHECK_ARG(mode == 0 || mode == 1, "vc_ln_modulate: mode must be 0 (modulate) or 1 (affine)");
  if (rows == 0) return 0;
  // Row ring (cp.async.bulk prefetch) for rows of >= 4 KB, register-resident kernel below that.  Measured on B200 at the c3
  // shape (profiles/rows_ab_r02.jsonl; in the step: profiles/bench_c3_r02c_{ring,regs}.json): LN+modulate 5.29 TB/s in the
  // step with the ring vs 4.04 with registers (alone 5.32 vs 5.04), norm3 affine 4.67 vs 3.84; at D = 1536 (3 KB rows) the
  // ring loses (3.95 vs 4.26 TB/s alone).  VC_ROW_KERNELS = ring | regs overrides; mode | 16 forces registers per call.
  static const int ring_env = []() {
    const char* e = getenv("VC_ROW_KERNELS");
    return !e ? -1 : !strcmp(e, "regs") ? 0 : !strcmp(e, "ring") ? 1 : -1;
  }();
  const bool use_ring = force_ring || (ring_env >= 0 ? ring_env == 1 : D >= 2048);
  if (D > 1024 && use_ring && !force_regs && reinterpret_cast<uintptr_t>(x) % 16 == 0) {
    int rcb = dispatch_vpt(D, [&](auto vpt) {
      constexpr int V = decltype(vpt)::value;
      const int smem = RING * D * 2;
      VC_ENSURE_DYN_SMEM((ln_modulate_ring_kernel<V>), smem, "vc_ln_modulate");
      ln_modulate_ring_kernel<V><<<grid_for_ring_rows(rows, V <= 5 ? 5 : 2), ROW_NT, smem, (cudaStream_t)stream>>>(
          (const __nv_bfloat16*)x, (__nv_bfloat16*)out, (const __nv_bfloat16*)shift_or_beta,
          (const __nv_bfloat16*)scale_or_gamma, mod_batch_stride, rows, rows_per_batch, D, ldx, ldo, eps, mode);
      return 0;
    });
    VC_CHECK_ARG(rcb == 0, "vc_ln_modulate: unsupported D=%d", D);
    VC_RETURN_LAUNCH_STATUS("vc_ln_modulate");
  }
  if (D > 1024) {
    int rcb = dispatch_vpt(D, [&](auto vpt) {
      constexpr int V = decltype(vpt)::value;
      ln_modulate_block_kernel<V><<<grid_for_block_rows(rows), ROW_NT, 0, (cudaStream_t)stream>>>(
          (const __nv_bfloat16*)x, (__nv_bfloat16*)out, (const __nv_bfloat16*)shift_or_beta,
          (const __nv_bfloat16*)scale_or_gamma, mod_batch_stride, rows, rows_per_batch, D, ldx, ldo, eps, mode);
      return 0;
    });
    VC_CHECK_ARG(rcb == 0, "vc_ln_modulate: unsupported D=%d", D);
    VC_RETURN_LAUNCH_STATUS("vc_ln_modulate");
  }
  int rc = dispatch_vpl(D, [&](auto vpl) {
    constexpr int V = decltype(vpl)::value;
    ln_modulate_kernel<V><<<grid_for_rows(rows, 8), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)x, (__nv_bfloat16*)out, (const __nv_bfloat16*)shift_or_beta,
        (const __nv_bfloat16*)scale_or_gamma, mod_batch_stride, rows, rows_per_batch, D, ldx, ldo, eps, mode);
    return 0;
  });
  VC_CHECK_ARG(rc == 0, "vc_ln_modulate: unsupported D=%d", D);
  VC_RETURN_LAUNCH_STATUS("vc_ln_modulate");
}

int vc_rmsnorm_rope_qk(void* qkv, long long ld, long long rows, long long rows_per_batch, long long pos_offset, int D,
                       int n_mats, const void* gain_q, const void* gain_k, const void* rope_table, int grid_f,
                       int grid_h, int grid_w, int nf, int nh, int nw, float eps, int apply_rope, void* stream) {
  const bool force_regs = (apply_rope & 16) != 0;  // schedule bits as in vc_ln_modulate: 16 = registers, 32 = row ring
  const bool force_ring = (apply_rope & 32) != 0;
  apply_rope &= 1;
  VC_CHECK_ARG(n_mats == 1 || n_mats == 2, "vc_rmsnorm_rope_qk: n_mats must be 1 or 2");
  VC_CHECK_ARG(qkv && gain_q && (n_mats == 1 || gain_k), "vc_rmsnorm_rope_qk: null pointer");
  VC_CHECK_ARG(D > 0 && D % 8 == 0 && D <= 8192 && ld % 8 == 0 && ld >= (long long)n_mats * D, "vc_rmsnorm_rope_qk: bad D/ld");
  VC_CHECK_ARG(!apply_rope || (rope_table && nf > 0 && nh > 0 && nw > 0 && (nf + nh + nw) % 4 == 0 &&
                               D % (2 * (nf + nh + nw)) == 0),
               "vc_rmsnorm_rope_qk: bad rope table split");
  VC_CHECK_ARG(rows >= 0 && rows_per_batch > 0 && pos_offset >= 0, "vc_rmsnorm_rope_qk: bad rows");
  VC_CHECK_ARG(rows < (1LL << 31) && pos_offset + rows_per_batch < (1LL << 31) &&
                   (!apply_rope || (long long)grid_f * grid_h * grid_w < (1LL << 31)),
               "vc_rmsnorm_rope_qk: rows, positions and the grid volume must be < 2^31");
  if (rows == 0) return 0;
  // The in-place RMSNorm(+RoPE) keeps the register-resident kernel (it already prefetches the next segment into registers):
  // the ring variant is on par alone at D = 5120 (5.80 vs 5.76 TB/s) but slower in the step (4.29 vs 4.77 TB/s) and at
  // D = 1536; it runs only with VC_ROW_KERNELS=ring (kept for experiments, bit-identical).
  static const bool ring_env = []() { const char* e = getenv("VC_ROW_KERNELS"); return e && !strcmp(e, "ring"); }();
  const bool use_ring = ring_env || force_ring;
  if (D > 1024 && use_ring && !force_regs && (!apply_rope || (ROW_NT * 8) % (2 * (nf + nh + nw)) == 0) &&
      reinterpret_cast<uintptr_t>(qkv) % 16 == 0) {
    int rcb = dispatch_vpt(D, [&](auto vpt) {
      constexpr int V = decltype(vpt)::value;
      const int smem = RING * D * 2;
      VC_ENSURE_DYN_SMEM((rmsnorm_rope_ring_kernel<V>), smem, "vc_rmsnorm_rope_qk");
      rmsnorm_rope_ring_kernel<V><<<grid_for_ring_rows(rows, V <= 5 ? 5 : 4), ROW_NT, smem, (cudaStream_t)stream>>>(
          (__nv_bfloat16*)qkv, ld, rows, rows_per_batch, pos_offset, D, (const __nv_bfloat16*)gain_q,
          (const __nv_bfloat16*)gain_k, (const float2*)rope_table, grid_f, grid_h, grid_w, nf, nh, nw, eps, apply_rope,
          n_mats);
      return 0;
    });
    VC_CHECK_ARG(rcb == 0, "vc_rmsnorm_rope_qk: unsupported D=%d", D);
    VC_RETURN_LAUNCH_STATUS("vc_rmsnorm_rope_qk");
  }
  // Self-attention q|k with RoPE is the case vc_rmsnorm_rope_qkv_scatter mirrors under sequence parallelism; that kernel is
  // block-per-row for every D, so this one is too (same fp32 summation order => the sequence-parallel forward is bit-identical
  // to the single-GPU one also at D <= 1024, where the warp-per-row kernel would otherwise round the sum of squares differently
  // in about one row per million: tools/sp_diag.py).
  const bool mirrored = n_mats == 2 && apply_rope;
  if ((D > 1024 || mirrored) && (!apply_rope || (ROW_NT * 8) % (2 * (nf + nh + nw)) == 0)) {
    int rcb = dispatch_vpt(D, [&](auto vpt) {
      constexpr int V = decltype(vpt)::value;
      rmsnorm_rope_block_kernel<V><<<grid_for_block_rows(rows), ROW_NT, 0, (cudaStream_t)stream>>>(
          (__nv_bfloat16*)qkv, ld, rows, rows_per_batch, pos_offset, D, (const __nv_bfloat16*)gain_q,
          (const __nv_bfloat16*)gain_k, (const float2*)rope_table, grid_f, grid_h, grid_w, nf, nh, nw, eps, apply_rope,
          n_mats);
      return 0;
    });
    VC_CHECK_ARG(rcb == 0, "vc_rmsnorm_rope_qk: unsupported D=%d", D);
    VC_RETURN_LAUNCH_STATUS("vc_rmsnorm_rope_qk");
  }
  int rc = dispatch_vpl(D, [&](auto vpl) {
    constexpr int V = decltype(vpl)::value;
    rmsnorm_rope_kernel<V><<<grid_for_rows(rows, 8), 256, 0, (cudaStream_t)stream>>>(
        (__nv_bfloat16*)qkv, ld, rows, rows_per_batch, pos_offset, D, (const __nv_bfloat16*)gain_q,
        (const __nv_bfloat16*)gain_k, (const float2*)rope_table, grid_f, grid_h, grid_w, nf, nh, nw, eps, apply_rope, n_mats);
    return 0;
  });
  VC_CHECK_ARG(rc == 0, "vc_rmsnorm_rope_qk: unsupported D=%d", D);
  VC_RETURN_LAUNCH_STATUS("vc_rmsnorm_rope_qk");
}

int vc_rmsnorm_rope_qkv_scatter(const void* qkv, long long ld, long long rows, long long rows_per_batch,
                                long long pos_offset, int D, const void* gain_q, const void* gain_k,
                                const void* rope_table, int grid_f, int grid_h, int grid_w, int nf, int nh, int nw,
                                float eps, void* const* peer_recv, int n_peers, int heads_per_peer, long long l_total,
                                int which_mask, void* stream) {
  VC_CHECK_ARG(qkv && gain_q && gain_k && rope_table && peer_recv, "vc_rmsnorm_rope_qkv_scatter: null pointer");
  // which_mask: bits 0-2 select q / k / v; bit 3 (8) = bulk-store variant (stores issued by the TMA engine); bits 8-15 =
  // cap on resident CTAs per SM for this launch (0 = default), used when the launch runs under a GEMM on another stream
  const bool bulk = (which_mask & 8) != 0;
  const int cta_cap = (which_mask >> 8) & 0xff;
  which_mask &= 7;
  VC_CHECK_ARG(which_mask >= 1, "vc_rmsnorm_rope_qkv_scatter: which_mask selects no matrix");
  VC_CHECK_ARG(n_peers >= 1 && n_peers <= 8, "vc_rmsnorm_rope_qkv_scatter: n_peers=%d must be in 1..8", n_peers);
  VC_CHECK_ARG(nf > 0 && nh > 0 && nw > 0 && (nf + nh + nw) % 4 == 0, "vc_rmsnorm_rope_qkv_scatter: bad rope split");
  const int hd = 2 * (nf + nh + nw);
  VC_CHECK_ARG(D % hd == 0 && (ROW_NT * 8) % hd == 0 && D <= 8192 && ld % 8 == 0 && ld >= 3LL * D,
               "vc_rmsnorm_rope_qkv_scatter: unsupported D=%d / head_dim=%d / ld", D, hd);
  VC_CHECK_ARG(heads_per_peer > 0 && (long long)heads_per_peer * n_peers == D / hd,
               "vc_rmsnorm_rope_qkv_scatter: heads_per_peer * n_peers must equal the head count");
  VC_CHECK_ARG(rows >= 0 && rows_per_batch > 0 && rows % rows_per_batch == 0 && pos_offset >= 0 &&
                   pos_offset + rows_per_batch <= l_total,
               "vc_rmsnorm_rope_qkv_scatter: bad rows / positions");
  VC_CHECK_ARG(rows < (1LL << 31) && l_total < (1LL << 31) && (long long)grid_f * grid_h * grid_w < (1LL << 31),
               "vc_rmsnorm_rope_qkv_scatter: rows, positions and the grid volume must be < 2^31");
  if (rows == 0) return 0;
  PeerPtrs pp;
  for (int i = 0; i < 8; ++i) pp.p[i] = (__nv_bfloat16*)(i < n_peers ? peer_recv[i] : nullptr);
  for (int i = 0; i < n_peers; ++i) VC_CHECK_ARG(pp.p[i] && (uintptr_t)pp.p[i] % 16 == 0, "vc_rmsnorm_rope_qkv_scatter: bad peer pointer %d", i);
  int grid = grid_for_block_rows(rows);
  if (cta_cap > 0 && grid > cta_cap * vc_host::sm_count()) grid = cta_cap * vc_host::sm_count();
  if (bulk) {
    int rcb = dispatch_vpt(D, [&](auto vpt) {
      constexpr int V = decltype(vpt)::value;
      const int smem = 2 * D * 2;
      VC_ENSURE_DYN_SMEM((rmsnorm_rope_scatter_bulk_kernel<V>), smem, "vc_rmsnorm_rope_qkv_scatter");
      const int cap8 = 8 * vc_host::sm_count();
      rmsnorm_rope_scatter_bulk_kernel<V><<<grid < cap8 ? grid : cap8, ROW_NT, smem, (cudaStream_t)stream>>>(
          (const __nv_bfloat16*)qkv, ld, rows, rows_per_batch, pos_offset, D, (const __nv_bfloat16*)gain_q,
          (const __nv_bfloat16*)gain_k, (const float2*)rope_table, grid_f, grid_h, grid_w, nf, nh, nw, eps, pp, n_peers,
          heads_per_peer, l_total, (int)(rows / rows_per_batch), which_mask);
      return 0;
    });
    VC_CHECK_ARG(rcb == 0, "vc_rmsnorm_rope_qkv_scatter: unsupported D=%d", D);
    VC_RETURN_LAUNCH_STATUS("vc_rmsnorm_rope_qkv_scatter");
  }
  int rc = dispatch_vpt(D, [&](auto vpt) {
    constexpr int V = decltype(vpt)::value;
    rmsnorm_rope_scatter_kernel<V><<<grid, ROW_NT, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)qkv, ld, rows, rows_per_batch, pos_offset, D, (const __nv_bfloat16*)gain_q,
        (const __nv_bfloat16*)gain_k, (const float2*)rope_table, grid_f, grid_h, grid_w, nf, nh, nw, eps, pp,
        heads_per_peer, l_total, (int)(rows / rows_per_batch), which_mask);
    return 0;
  });
  VC_CHECK_ARG(rc == 0, "vc_rmsnorm_rope_qkv_scatter: unsupported D=%d", D);
  VC_RETURN_LAUNCH_STATUS("vc_rmsnorm_rope_qkv_scatter");
}

int vc_patchify(const void* in, void* out, int C, int F, int H, int W, long long ld_out, void* stream) {
  VC_CHECK_ARG(in && out && C > 0 && F > 0 && H > 0 && W > 0 && H % 2 == 0 && W % 2 == 0, "vc_patchify: bad shape");
  VC_CHECK_ARG(ld_out >= 4LL * C && ld_out % 4 == 0, "vc_patchify: bad ld_out");
  long long total = (long long)F * (H / 2) * (W / 2) * C;
  int grid = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  patchify_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)in, (__nv_bfloat16*)out, C, F, H, W, ld_out);
  VC_RETURN_LAUNCH_STATUS("vc_patchify");
}

int vc_unpatchify(const void* y, long long ldy, void* out, int Co, int gf, int gh, int gw, void* stream) {
  VC_CHECK_ARG(y && out && Co > 0 && gf > 0 && gh > 0 && gw > 0 && ldy >= 4LL * Co, "vc_unpatchify: bad shape");
  long long total = (long long)Co * gf * gh * gw * 4;
  int grid = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  unpatchify_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)y, (__nv_bfloat16*)out, Co, gf, gh, gw, ldy);
  VC_RETURN_LAUNCH_STATUS("vc_unpatchify");
}

}  // extern "C"
