// Non-causal flash attention forward, head_dim 128, bf16 in/out, fp32 softmax state, on tcgen05 tensor cores with
// TMEM-resident S / O accumulators (SURVEY.md §2.2 K6 self-attention, K7 cross-attention).
// Replaces videox_fun attention() at wan_transformer3d.py:394-399 and :425-430:  softmax(Q K^T / sqrt(d)) V,
// keys beyond Lk masked (TMA zero-fills the tail; the last key tile is masked to -inf before the row max).
//
// One CTA = one (batch, head, 256 query rows) = two 128-row Q tiles that ping-pong on the tensor pipe:
//   warp 0      TMA producer: Q once, then K_0,V_0,K_1,V_1,... through an mbarrier ring of 32 KB slots
//   warp 1      one elected thread issues every tcgen05.mma:  S_q = Q_q K_j^T   (M=128,N=128,K=128)
//                                                              O_q += P_q V_j   (M=128,N=128,K=128, V MN-major)
//   warp 2      TMEM allocator (512 columns: S0 | S1 | O0 | O1)
//   warps 4-7   softmax of Q tile 0 (thread = query row = TMEM lane), warps 8-11 the same for Q tile 1:
//               tcgen05.ld S row -> running max (lazy rescale, threshold 2^8) -> exp2 -> P (bf16)
//               P goes to shared memory (kPInTmem=false, SS MMA) or back into TMEM over S (kPInTmem=true, TS MMA)
// Issue order  S0_0 S1_0 | PV0_0 S0_1 | PV1_0 S1_1 | PV0_1 S0_2 | ...  so that the softmax of one Q tile runs while
// the tensor pipe works on the other.  tcgen05 ops of one thread retire in order, so "S_q,j+1 done" implies
// "PV_q,j done": that single s_full barrier also guards the O_q rescale and the reuse of the P_q buffer.
#include "common.cuh"

namespace {

constexpr int HD = 128;        // head dim
constexpr int BQ = 128;        // rows per Q tile (2 tiles per CTA)
constexpr int BKV = 128;       // keys per tile
constexpr int HALF_BYTES = 128 * 128;          // 128 rows x 64 bf16 (one 128B-swizzle column block) = 16 KB
constexpr int TILE_BYTES = 2 * HALF_BYTES;     // 32 KB
constexpr int NUM_THREADS = 384;
constexpr float RESCALE_THRESHOLD = 8.0f;      // log2 units

template <bool kPInTmem>
struct Cfg {
  static constexpr int NSLOT = kPInTmem ? 5 : 3;
  static constexpr int P_BYTES = kPInTmem ? 0 : 2 * TILE_BYTES;
  static constexpr int SMEM_BYTES = 2 * TILE_BYTES + NSLOT * TILE_BYTES + P_BYTES + 256 + 1024;
};

struct AttnParams {
  __nv_bfloat16* O;
  long long o_stride_b, o_stride_l, o_stride_h;  // elements
  int Lq, Lk;
  float scale_log2;  // softmax_scale * log2(e)
  // Ulysses output scatter (variants >= 2): query row t belongs to peer t / rows_per_peer and is stored at row
  // t % rows_per_peer of o_peer[that peer] (same strides as O).  n_peers == 0: plain store to O.
  __nv_bfloat16* o_peer[8];
  int n_peers;
  int rows_per_peer;
};

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;\n" : "=f"(y) : "f"(x));
  return y;
}

// packed fp32x2 arithmetic (sm_100 FFMA2 / FADD2): one issue slot for two lanes of the softmax row
__device__ __forceinline__ void ffma2(float& d0, float& d1, float a0, float a1, float b0, float b1, float c0, float c1) {
  asm("{\n\t.reg .b64 ra, rb, rc, rd;\n\t"
      "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\tmov.b64 rc, {%6, %7};\n\t"
      "fma.rn.ftz.f32x2 rd, ra, rb, rc;\n\t"
      "mov.b64 {%0, %1}, rd;\n\t}\n"
      : "=f"(d0), "=f"(d1)
      : "f"(a0), "f"(a1), "f"(b0), "f"(b1), "f"(c0), "f"(c1));
}
__device__ __forceinline__ void fadd2(float& d0, float& d1, float a0, float a1, float b0, float b1) {
  asm("{\n\t.reg .b64 ra, rb, rd;\n\t"
      "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\t"
      "add.rn.ftz.f32x2 rd, ra, rb;\n\t"
      "mov.b64 {%0, %1}, rd;\n\t}\n"
      : "=f"(d0), "=f"(d1)
      : "f"(a0), "f"(a1), "f"(b0), "f"(b1));
}
// 2^x for two values on the FMA/ALU pipes instead of the MUFU (Cody-Waite split + degree-3 minimax polynomial for
// 2^f on [0,1), rel. error < 1e-4: below the bf16 rounding of P).  x <= ~8 here; clamped at -126.
__device__ __forceinline__ void exp2_poly2(float& r0, float& r1, float x0, float x1) {
  constexpr float MAGIC = 12582912.0f;  // 1.5 * 2^23: x + MAGIC (round toward -inf) holds floor(x) in the low mantissa
  x0 = fmaxf(x0, -126.0f);
  x1 = fmaxf(x1, -126.0f);
  float t0, t1;
  asm("add.rm.ftz.f32 %0, %1, %2;" : "=f"(t0) : "f"(x0), "f"(MAGIC));
  asm("add.rm.ftz.f32 %0, %1, %2;" : "=f"(t1) : "f"(x1), "f"(MAGIC));
  float fl0, fl1, f0, f1;
  fadd2(fl0, fl1, t0, t1, -MAGIC, -MAGIC);   // floor(x)
  fadd2(f0, f1, x0, x1, -fl0, -fl1);         // fractional part in [0, 1)
  float p0, p1;
  ffma2(p0, p1, f0, f1, 0.07802334f, 0.07802334f, 0.22606642f, 0.22606642f);
  ffma2(p0, p1, p0, p1, f0, f1, 0.69583512f, 0.69583512f);
  ffma2(p0, p1, p0, p1, f0, f1, 0.99992492f, 0.99992492f);
  r0 = __int_as_float(__float_as_int(p0) + (__float_as_int(t0) << 23));
  r1 = __int_as_float(__float_as_int(p1) + (__float_as_int(t1) << 23));
}

template <bool kPInTmem>
__global__ void __launch_bounds__(NUM_THREADS, 1)
flash_attn_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                  const __grid_constant__ CUtensorMap tmV, const AttnParams p) {
  using C = Cfg<kPInTmem>;
  constexpr int NSLOT = C::NSLOT;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                                   // 2 tiles
  uint8_t* sKV = smem + 2 * TILE_BYTES;                 // NSLOT tiles
  uint8_t* sP = sKV + NSLOT * TILE_BYTES;               // 2 tiles (SS variant only)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + C::P_BYTES);
  uint64_t* q_full = bars;            // 1
  uint64_t* kv_full = bars + 1;       // NSLOT
  uint64_t* kv_empty = kv_full + NSLOT;
  uint64_t* s_full = kv_empty + NSLOT;  // 2
  uint64_t* p_ready = s_full + 2;       // 2
  uint64_t* o_full = p_ready + 2;       // 2
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int q_blk = blockIdx.x, head = blockIdx.y, batch = blockIdx.z;
  const int nk = (p.Lk + BKV - 1) / BKV;

  if (threadIdx.x == 0) {
    vc::mbar_init(q_full, 1);
    for (int i = 0; i < NSLOT; ++i) {
      vc::mbar_init(&kv_full[i], 1);
      vc::mbar_init(&kv_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      vc::mbar_init(&s_full[i], 1);
      vc::mbar_init(&p_ready[i], 128);
      vc::mbar_init(&o_full[i], 1);
    }
    vc::fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    vc::tma_prefetch_desc(&tmQ);
    vc::tma_prefetch_desc(&tmK);
    vc::tma_prefetch_desc(&tmV);
  }
  if (warp == 2) vc::tmem_alloc(tmem_slot, 512);
  vc::tc_fence_before();
  __syncthreads();
  vc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 4) {
    vc::reg_dealloc<56>();
    if (warp == 0) {
      // ---------------------------------------------------------------- TMA producer
      if (vc::elect_one()) {
        vc::mbar_expect_tx(q_full, 2 * TILE_BYTES);
#pragma unroll
        for (int qt = 0; qt < 2; ++qt)
#pragma unroll
          for (int h = 0; h < 2; ++h)
            vc::tma_load_4d(sQ + qt * TILE_BYTES + h * HALF_BYTES, &tmQ, q_full, h * 64, (q_blk * 2 + qt) * BQ, head,
                            batch);
        for (int idx = 0; idx < 2 * nk; ++idx) {
          const int slot = idx % NSLOT;
          const uint32_t ph = (idx / NSLOT) & 1;
          vc::mbar_wait(&kv_empty[slot], ph ^ 1);
          vc::mbar_expect_tx(&kv_full[slot], TILE_BYTES);
          const CUtensorMap* tm = (idx & 1) ? &tmV : &tmK;
          const int row0 = (idx >> 1) * BKV;
#pragma unroll
          for (int h = 0; h < 2; ++h)
            vc::tma_load_4d(sKV + slot * TILE_BYTES + h * HALF_BYTES, tm, &kv_full[slot], h * 64, row0, head, batch);
        }
      }
    } else if (warp == 1) {
      // ---------------------------------------------------------------- MMA issuer
      if (vc::elect_one()) {
        constexpr uint32_t idesc_s = vc::umma_idesc_bf16(BQ, BKV, 0, 0);  // Q (K-major) x K (K-major)
        constexpr uint32_t idesc_o = vc::umma_idesc_bf16(BQ, HD, 0, 1);   // P (K-major) x V (MN-major)
        auto kv_wait = [&](int idx) { vc::mbar_wait(&kv_full[idx % NSLOT], (idx / NSLOT) & 1); };
        auto kv_addr = [&](int idx) { return vc::smem_u32(sKV + (idx % NSLOT) * TILE_BYTES); };
        auto issue_s = [&](int q, int j) {
          const uint32_t qa = vc::smem_u32(sQ + q * TILE_BYTES);
          const uint32_t ka = kv_addr(2 * j);
#pragma unroll
          for (int k = 0; k < HD / 16; ++k) {
            const uint32_t off = (k >> 2) * HALF_BYTES + (k & 3) * 32;
            vc::umma_ss(tmem_base + q * 128, vc::umma_desc_sw128(qa + off, 16, 1024),
                        vc::umma_desc_sw128(ka + off, 16, 1024), idesc_s, k != 0 ? 1u : 0u);
          }
          vc::umma_commit(&s_full[q]);
        };
        auto issue_pv = [&](int q, int j) {
          const uint32_t va = kv_addr(2 * j + 1);
          const uint32_t d_tmem = tmem_base + 256 + q * 128;
#pragma unroll
          for (int k = 0; k < BKV / 16; ++k) {
            const uint64_t vdesc = vc::umma_desc_sw128(va + k * 2048, HALF_BYTES, 1024);
            const uint32_t accum = (j != 0 || k != 0) ? 1u : 0u;
            if (kPInTmem) {
              vc::umma_ts(d_tmem, tmem_base + q * 128 + k * 8, vdesc, idesc_o, accum);
            } else {
              const uint32_t pa = vc::smem_u32(sP + q * TILE_BYTES) + (k >> 2) * HALF_BYTES + (k & 3) * 32;
              vc::umma_ss(d_tmem, vc::umma_desc_sw128(pa, 16, 1024), vdesc, idesc_o, accum);
            }
          }
        };
        vc::mbar_wait(q_full, 0);
        kv_wait(0);
        vc::tc_fence_after();
        issue_s(0, 0);
        issue_s(1, 0);
        vc::umma_commit(&kv_empty[0]);
        for (int j = 0; j < nk; ++j) {
          kv_wait(2 * j + 1);  // V_j
          for (int q = 0; q < 2; ++q) {
            vc::mbar_wait(&p_ready[q], j & 1);
            vc::tc_fence_after();
            issue_pv(q, j);
            if (q == 1) vc::umma_commit(&kv_empty[(2 * j + 1) % NSLOT]);
            if (j + 1 < nk) {
              if (q == 0) {
                kv_wait(2 * j + 2);  // K_{j+1}
                vc::tc_fence_after();
              }
              issue_s(q, j + 1);
              if (q == 1) vc::umma_commit(&kv_empty[(2 * j + 2) % NSLOT]);
            } else {
              vc::umma_commit(&o_full[q]);
            }
          }
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax + correction + epilogue
    vc::reg_alloc<208>();
    const int q = (warp - 4) >> 2;          // Q tile of this warpgroup
    const int quarter = warp & 3;           // TMEM lane quarter this warp may touch
    const int row = quarter * 32 + lane;    // row inside the Q tile == TMEM lane
    const uint32_t lane_bits = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t s_addr = tmem_base + q * 128 + lane_bits;
    const uint32_t o_addr = tmem_base + 256 + q * 128 + lane_bits;
    const float sl2 = p.scale_log2;
    float m_run = -INFINITY;  // running (possibly stale) row max, raw score units
    float l_run = 0.f;
    uint8_t* p_row = sP + q * TILE_BYTES + row * 128;

    for (int j = 0; j < nk; ++j) {
      vc::mbar_wait(&s_full[q], j & 1);
      vc::tc_fence_after();
      uint32_t s[128];
#pragma unroll
      for (int c = 0; c < 4; ++c) vc::tmem_ld32(s_addr + c * 32, s + c * 32);
      vc::tmem_ld_wait();
      if (j == nk - 1) {
        const int valid = p.Lk - j * BKV;
        if (valid < BKV) {
#pragma unroll
          for (int i = 0; i < 128; ++i)
            if (i >= valid) s[i] = __float_as_uint(-INFINITY);
        }
      }
      // 8 independent max chains (a single 127-deep dependent FMNMX chain costs ~500 cycles of pure latency)
      float mx[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) mx[i] = __uint_as_float(s[i]);
#pragma unroll
      for (int i = 8; i < 128; ++i) mx[i & 7] = fmaxf(mx[i & 7], __uint_as_float(s[i]));
      const float mt = fmaxf(fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])), fmaxf(fmaxf(mx[4], mx[5]), fmaxf(mx[6], mx[7])));
      const float m_new = fmaxf(m_run, mt);
      const bool need = (m_new - m_run) * sl2 > RESCALE_THRESHOLD;  // true on the first tile (m_run = -inf)
      if (__any_sync(0xffffffffu, need)) {
        const float alpha = fast_exp2((m_run - m_new) * sl2);
        l_run *= alpha;
        m_run = m_new;
        if (j > 0) {
          // O_q is idle here: S_q,j done implies PV_q,j-1 done, and PV_q,j waits for our p_ready arrival
#pragma unroll 1
          for (int c = 0; c < 4; ++c) {
            uint32_t o[32];
            vc::tmem_ld32(o_addr + c * 32, o);
            vc::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            vc::tmem_st32(o_addr + c * 32, o);
          }
          vc::tmem_st_wait();
        }
      }
      const float mb = m_run * sl2;
      float sum = 0.f;
      if (kPInTmem) {
        float sum1 = 0.f;
#pragma unroll
        for (int i = 0; i < 64; ++i) {
          const float p0 = fast_exp2(fmaf(__uint_as_float(s[2 * i]), sl2, -mb));
          const float p1 = fast_exp2(fmaf(__uint_as_float(s[2 * i + 1]), sl2, -mb));
          sum += p0;
          sum1 += p1;
          s[i] = vc::pack_bf16x2(p0, p1);
        }
        sum += sum1;
        vc::tmem_st32(s_addr, s);
        vc::tmem_st32(s_addr + 32, s + 32);
        vc::tmem_st_wait();
        l_run += sum;
        vc::tc_fence_before();
      } else {
#pragma unroll
        for (int ch = 0; ch < 16; ++ch) {  // 8 keys = one 16-byte chunk of the swizzled K-major P tile
          float pv[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            pv[i] = fast_exp2(fmaf(__uint_as_float(s[ch * 8 + i]), sl2, -mb));
            sum += pv[i];
          }
          uint4 u;
          u.x = vc::pack_bf16x2(pv[0], pv[1]);
          u.y = vc::pack_bf16x2(pv[2], pv[3]);
          u.z = vc::pack_bf16x2(pv[4], pv[5]);
          u.w = vc::pack_bf16x2(pv[6], pv[7]);
          const int half = ch >> 3, chunk = ch & 7;
          *reinterpret_cast<uint4*>(p_row + half * HALF_BYTES + ((chunk ^ (row & 7)) << 4)) = u;
        }
        l_run += sum;
        vc::fence_proxy_async_smem();
      }
      vc::mbar_arrive(&p_ready[q]);
    }

    // epilogue: O_q / l -> bf16 -> global
    vc::mbar_wait(&o_full[q], 0);
    vc::tc_fence_after();
    const float inv_l = 1.0f / l_run;
    const int q_row = (q_blk * 2 + q) * BQ + row;
    __nv_bfloat16* orow = p.O + (long long)batch * p.o_stride_b + (long long)q_row * p.o_stride_l +
                          (long long)head * p.o_stride_h;
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      uint32_t o[32];
      vc::tmem_ld32(o_addr + c * 32, o);
      vc::tmem_ld_wait();
      if (q_row < p.Lq) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 u;
          u.x = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 0]) * inv_l, __uint_as_float(o[g * 8 + 1]) * inv_l);
          u.y = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 2]) * inv_l, __uint_as_float(o[g * 8 + 3]) * inv_l);
          u.z = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 4]) * inv_l, __uint_as_float(o[g * 8 + 5]) * inv_l);
          u.w = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 6]) * inv_l, __uint_as_float(o[g * 8 + 7]) * inv_l);
          *reinterpret_cast<uint4*>(orow + c * 32 + g * 8) = u;
        }
      }
    }
  }

  vc::tc_fence_before();
  __syncthreads();
  if (warp == 2) vc::tmem_dealloc(tmem_base, 512);
}

int make_qkv_map(CUtensorMap* tm, const void* base, int B, int H, int L, long long sb, long long sl, long long sh,
                 uint32_t box_rows = 128) {
  uint64_t dims[4] = {(uint64_t)HD, (uint64_t)L, (uint64_t)H, (uint64_t)B};
  uint64_t strides[3] = {(uint64_t)sl * 2, (uint64_t)sh * 2, (uint64_t)sb * 2};
  uint32_t box[4] = {64, box_rows, 1, 1};
  return vc_host::make_tmap_bf16(tm, base, 4, dims, strides, box);
}

template <bool kPInTmem>
int launch(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AttnParams& p, dim3 grid,
           cudaStream_t s) {
  constexpr int smem = Cfg<kPInTmem>::SMEM_BYTES;
  VC_ENSURE_DYN_SMEM((flash_attn_kernel<kPInTmem>), smem, "vc_flash_attn_fwd");
  flash_attn_kernel<kPInTmem><<<grid, NUM_THREADS, smem, s>>>(tq, tk, tv, p);
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// Variant 2 ("shared S"): the S -> softmax -> PV -> S chain of one Q tile is cut.
// In variants 0/1 S_q,j+1 may only be issued after PV_q,j has consumed P_q,j (P aliases S), so each Q tile runs
// S(512 cyc) -> softmax(~1800) -> PV(512) strictly in series and the tensor pipe idles 36 % of the time (ncu, round 1:
// the softmax warps spend 45 % of their samples waiting for S).  Here TMEM is laid out
//     SA (128 cols, ONE S buffer shared by both Q tiles) | P0 (64) | P1 (64) | O0 (128) | O1 (128)   = 512 columns
// A softmax warpgroup copies S into registers as soon as it is complete and releases SA at once (sa_free), so the
// next S (of the other Q tile) is computed while both warpgroups are still busy with exp2; P_q lives in its own
// columns, so PV_q,j never blocks an S.  S_q,j+1 is therefore ready long before softmax_q,j ends: the softmax warps
// never wait for the tensor pipe and the two pipes (tensor, MUFU) overlap instead of alternating.
// MMA issue order (one thread):  S0_0 S1_0 S0_1 | PV0_0 S1_1 PV1_0 S0_2 | PV0_1 S1_2 PV1_1 S0_3 | ...
// K/V tiles are loaded in first-use order  K0 K1 V0 K2 V1 K3 V2 ...  (ring of 5 x 32 KB slots, freed in the same order).
// ---------------------------------------------------------------------------------------------------------------
constexpr int SA_NSLOT = 5;
constexpr int SA_SMEM_BYTES = 2 * TILE_BYTES + SA_NSLOT * TILE_BYTES + 256 + 1024;

// kTuned: PV_0,j is issued in two halves around S_1,j+1 (so the other Q tile's S does not queue behind a full PV),
// the softmax row uses packed fp32x2 FMA/ADD, and kPoly of every 128 exponentials run on the FMA pipe (exp2_poly2).
template <bool kTuned, int kPoly>
__global__ void __launch_bounds__(NUM_THREADS, 1)
flash_attn_sa_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                     const __grid_constant__ CUtensorMap tmV, const AttnParams p) {
  constexpr int NSLOT = SA_NSLOT;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                    // 2 tiles
  uint8_t* sKV = smem + 2 * TILE_BYTES;  // NSLOT tiles
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + NSLOT * TILE_BYTES);
  uint64_t* q_full = bars;               // 1
  uint64_t* kv_full = bars + 1;          // NSLOT
  uint64_t* kv_empty = kv_full + NSLOT;  // NSLOT
  uint64_t* s_full = kv_empty + NSLOT;   // 2  (MMA -> softmax q: S_q is in SA)
  uint64_t* sa_free = s_full + 2;        // 1  (softmax -> MMA: SA copied to registers)
  uint64_t* p_ready = sa_free + 1;       // 2  (softmax q -> MMA: P_q written)
  uint64_t* pv_done = p_ready + 2;       // 2  (MMA -> softmax q: PV_q,j retired: P_q reusable, O_q idle)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(pv_done + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int q_blk = blockIdx.x, head = blockIdx.y, batch = blockIdx.z;
  const int nk = (p.Lk + BKV - 1) / BKV;

  if (threadIdx.x == 0) {
    vc::mbar_init(q_full, 1);
    for (int i = 0; i < NSLOT; ++i) {
      vc::mbar_init(&kv_full[i], 1);
      vc::mbar_init(&kv_empty[i], 1);
    }
    vc::mbar_init(sa_free, 128);
    for (int i = 0; i < 2; ++i) {
      vc::mbar_init(&s_full[i], 1);
      vc::mbar_init(&p_ready[i], 128);
      vc::mbar_init(&pv_done[i], 1);
    }
    vc::fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    vc::tma_prefetch_desc(&tmQ);
    vc::tma_prefetch_desc(&tmK);
    vc::tma_prefetch_desc(&tmV);
  }
  if (warp == 2) vc::tmem_alloc(tmem_slot, 512);
  vc::tc_fence_before();
  __syncthreads();
  vc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  constexpr uint32_t COL_SA = 0, COL_P = 128, COL_O = 256;

  // position of a tile in the load / use / free sequence K0 K1 V0 K2 V1 K3 V2 ... K_{nk-1} V_{nk-2} V_{nk-1}
  auto pos_k = [&](int j) { return j == 0 ? 0 : 2 * j - 1; };
  auto pos_v = [&](int j) { return j == nk - 1 ? 2 * nk - 1 : 2 * j + 2; };

  if (warp < 4) {
    vc::reg_dealloc<56>();
    if (warp == 0) {
      // ---------------------------------------------------------------- TMA producer
      if (vc::elect_one()) {
        vc::mbar_expect_tx(q_full, 2 * TILE_BYTES);
#pragma unroll
        for (int qt = 0; qt < 2; ++qt)
#pragma unroll
          for (int h = 0; h < 2; ++h)
            vc::tma_load_4d(sQ + qt * TILE_BYTES + h * HALF_BYTES, &tmQ, q_full, h * 64, (q_blk * 2 + qt) * BQ, head,
                            batch);
        auto load = [&](int pos, const CUtensorMap* tm, int j) {
          const int slot = pos % NSLOT;
          const uint32_t ph = (pos / NSLOT) & 1;
          vc::mbar_wait(&kv_empty[slot], ph ^ 1);
          vc::mbar_expect_tx(&kv_full[slot], TILE_BYTES);
#pragma unroll
          for (int h = 0; h < 2; ++h)
            vc::tma_load_4d(sKV + slot * TILE_BYTES + h * HALF_BYTES, tm, &kv_full[slot], h * 64, j * BKV, head, batch);
        };
        load(0, &tmK, 0);
        for (int j = 0; j < nk; ++j) {
          if (j + 1 < nk) load(pos_k(j + 1), &tmK, j + 1);
          load(pos_v(j), &tmV, j);
        }
      }
    } else if (warp == 1) {
      // ---------------------------------------------------------------- MMA issuer
      if (vc::elect_one()) {
        constexpr uint32_t idesc_s = vc::umma_idesc_bf16(BQ, BKV, 0, 0);  // Q (K-major) x K (K-major)
        constexpr uint32_t idesc_o = vc::umma_idesc_bf16(BQ, HD, 0, 1);   // P (TMEM)    x V (MN-major)
        auto kv_wait = [&](int pos) { vc::mbar_wait(&kv_full[pos % NSLOT], (pos / NSLOT) & 1); };
        auto kv_addr = [&](int pos) { return vc::smem_u32(sKV + (pos % NSLOT) * TILE_BYTES); };
        auto kv_release = [&](int pos) { vc::umma_commit(&kv_empty[pos % NSLOT]); };
        uint32_t sa_uses = 0;  // completed sa_free phases consumed so far
        auto wait_sa_free = [&]() {
          vc::mbar_wait(sa_free, sa_uses & 1);
          ++sa_uses;
          vc::tc_fence_after();
        };
        auto issue_s = [&](int q, int j) {
          const uint32_t qa = vc::smem_u32(sQ + q * TILE_BYTES);
          const uint32_t ka = kv_addr(pos_k(j));
#pragma unroll
          for (int k = 0; k < HD / 16; ++k) {
            const uint32_t off = (k >> 2) * HALF_BYTES + (k & 3) * 32;
            vc::umma_ss(tmem_base + COL_SA, vc::umma_desc_sw128(qa + off, 16, 1024),
                        vc::umma_desc_sw128(ka + off, 16, 1024), idesc_s, k != 0 ? 1u : 0u);
          }
          vc::umma_commit(&s_full[q]);
        };
        auto issue_pv_part = [&](int q, int j, int k0, int k1) {
          const uint32_t va = kv_addr(pos_v(j));
          const uint32_t d_tmem = tmem_base + COL_O + q * 128;
          for (int k = k0; k < k1; ++k) {
            const uint64_t vdesc = vc::umma_desc_sw128(va + k * 2048, HALF_BYTES, 1024);
            vc::umma_ts(d_tmem, tmem_base + COL_P + q * 64 + k * 8, vdesc, idesc_o, (j != 0 || k != 0) ? 1u : 0u);
          }
          if (k1 == BKV / 16) vc::umma_commit(&pv_done[q]);
        };
        auto issue_pv = [&](int q, int j) { issue_pv_part(q, j, 0, BKV / 16); };
        vc::mbar_wait(q_full, 0);
        kv_wait(pos_k(0));
        vc::tc_fence_after();
        issue_s(0, 0);
        wait_sa_free();
        issue_s(1, 0);
        kv_release(pos_k(0));
        if (nk > 1) {
          wait_sa_free();
          kv_wait(pos_k(1));
          vc::tc_fence_after();
          issue_s(0, 1);
        }
        for (int j = 0; j < nk; ++j) {
          vc::mbar_wait(&p_ready[0], j & 1);
          kv_wait(pos_v(j));
          vc::tc_fence_after();
          if (kTuned && j + 1 < nk) {
            issue_pv_part(0, j, 0, BKV / 32);
            wait_sa_free();
            issue_s(1, j + 1);
            kv_release(pos_k(j + 1));
            issue_pv_part(0, j, BKV / 32, BKV / 16);
          } else {
            issue_pv(0, j);
            if (j + 1 < nk) {
              wait_sa_free();
              issue_s(1, j + 1);
              kv_release(pos_k(j + 1));
            }
          }
          vc::mbar_wait(&p_ready[1], j & 1);
          vc::tc_fence_after();
          if (kTuned && j + 2 < nk) {
            issue_pv_part(1, j, 0, BKV / 32);
            wait_sa_free();
            kv_wait(pos_k(j + 2));
            vc::tc_fence_after();
            issue_s(0, j + 2);
            issue_pv_part(1, j, BKV / 32, BKV / 16);
            kv_release(pos_v(j));
          } else {
            issue_pv(1, j);
            kv_release(pos_v(j));
            if (j + 2 < nk) {
              wait_sa_free();
              kv_wait(pos_k(j + 2));
              vc::tc_fence_after();
              issue_s(0, j + 2);
            }
          }
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax + correction + epilogue
    vc::reg_alloc<208>();
    const int q = (warp - 4) >> 2;        // Q tile of this warpgroup
    const int quarter = warp & 3;         // TMEM lane quarter this warp may touch
    const int row = quarter * 32 + lane;  // row inside the Q tile == TMEM lane
    const uint32_t lane_bits = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t s_addr = tmem_base + COL_SA + lane_bits;
    const uint32_t p_addr = tmem_base + COL_P + q * 64 + lane_bits;
    const uint32_t o_addr = tmem_base + COL_O + q * 128 + lane_bits;
    const float sl2 = p.scale_log2;
    float m_run = -INFINITY;  // running (possibly stale) row max, raw score units
    float l_run = 0.f;

    for (int j = 0; j < nk; ++j) {
      vc::mbar_wait(&s_full[q], j & 1);
      vc::tc_fence_after();
      uint32_t s[128];
#pragma unroll
      for (int c = 0; c < 4; ++c) vc::tmem_ld32(s_addr + c * 32, s + c * 32);
      vc::tmem_ld_wait();
      vc::tc_fence_before();
      vc::mbar_arrive(sa_free);  // SA may be overwritten by the other Q tile's S
      if (j == nk - 1) {
        const int valid = p.Lk - j * BKV;
        if (valid < BKV) {
#pragma unroll
          for (int i = 0; i < 128; ++i)
            if (i >= valid) s[i] = __float_as_uint(-INFINITY);
        }
      }
      float mx[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) mx[i] = __uint_as_float(s[i]);
#pragma unroll
      for (int i = 8; i < 128; ++i) mx[i & 7] = fmaxf(mx[i & 7], __uint_as_float(s[i]));
      const float mt = fmaxf(fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])), fmaxf(fmaxf(mx[4], mx[5]), fmaxf(mx[6], mx[7])));
      const float m_new = fmaxf(m_run, mt);
      const bool need = (m_new - m_run) * sl2 > RESCALE_THRESHOLD;  // true on the first tile (m_run = -inf)
      bool waited_pv = false;
      if (__any_sync(0xffffffffu, need)) {
        const float alpha = fast_exp2((m_run - m_new) * sl2);
        l_run *= alpha;
        m_run = m_new;
        if (j > 0) {
          vc::mbar_wait(&pv_done[q], (j - 1) & 1);  // O_q idle: PV_q,j-1 retired and PV_q,j waits for our p_ready
          vc::tc_fence_after();
          waited_pv = true;
#pragma unroll 1
          for (int c = 0; c < 4; ++c) {
            uint32_t o[32];
            vc::tmem_ld32(o_addr + c * 32, o);
            vc::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            vc::tmem_st32(o_addr + c * 32, o);
          }
          vc::tmem_st_wait();
        }
      }
      const float mb = m_run * sl2;
      float sum = 0.f, sum1 = 0.f;
      if (kTuned) {
        float sum2 = 0.f, sum3 = 0.f;
#pragma unroll
        for (int i = 0; i < 64; ++i) {
          float x0, x1, p0, p1;
          ffma2(x0, x1, __uint_as_float(s[2 * i]), __uint_as_float(s[2 * i + 1]), sl2, sl2, -mb, -mb);
          if (kPoly > 0 && (i % (64 / (kPoly > 0 ? kPoly / 2 : 1))) == 1) {  // kPoly/2 pairs, evenly spread over the row
            exp2_poly2(p0, p1, x0, x1);
          } else {
            p0 = fast_exp2(x0);
            p1 = fast_exp2(x1);
          }
          if (i & 1) fadd2(sum2, sum3, sum2, sum3, p0, p1);
          else fadd2(sum, sum1, sum, sum1, p0, p1);
          s[i] = vc::pack_bf16x2(p0, p1);
        }
        l_run += (sum + sum1) + (sum2 + sum3);
      } else {
#pragma unroll
        for (int i = 0; i < 64; ++i) {
          const float p0 = fast_exp2(fmaf(__uint_as_float(s[2 * i]), sl2, -mb));
          const float p1 = fast_exp2(fmaf(__uint_as_float(s[2 * i + 1]), sl2, -mb));
          sum += p0;
          sum1 += p1;
          s[i] = vc::pack_bf16x2(p0, p1);
        }
        l_run += sum + sum1;
      }
      if (j > 0 && !waited_pv) {
        vc::mbar_wait(&pv_done[q], (j - 1) & 1);  // P_q,j-1 consumed
        vc::tc_fence_after();
      }
      vc::tmem_st32(p_addr, s);
      vc::tmem_st32(p_addr + 32, s + 32);
      vc::tmem_st_wait();
      vc::tc_fence_before();
      vc::mbar_arrive(&p_ready[q]);
    }

    // epilogue: O_q / l -> bf16 -> global
    vc::mbar_wait(&pv_done[q], (nk - 1) & 1);
    vc::tc_fence_after();
    const float inv_l = 1.0f / l_run;
    const int q_row = (q_blk * 2 + q) * BQ + row;
    __nv_bfloat16* obase = p.O;
    int o_row = q_row;
    if (p.n_peers > 0 && q_row < p.Lq) {  // inverse all-to-all fused into the epilogue: store into the owner's buffer
      const int owner = q_row / p.rows_per_peer;
      obase = p.o_peer[owner];
      o_row = q_row - owner * p.rows_per_peer;
    }
    __nv_bfloat16* orow = obase + (long long)batch * p.o_stride_b + (long long)o_row * p.o_stride_l +
                          (long long)head * p.o_stride_h;
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      uint32_t o[32];
      vc::tmem_ld32(o_addr + c * 32, o);
      vc::tmem_ld_wait();
      if (q_row < p.Lq) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 u;
          u.x = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 0]) * inv_l, __uint_as_float(o[g * 8 + 1]) * inv_l);
          u.y = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 2]) * inv_l, __uint_as_float(o[g * 8 + 3]) * inv_l);
          u.z = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 4]) * inv_l, __uint_as_float(o[g * 8 + 5]) * inv_l);
          u.w = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 6]) * inv_l, __uint_as_float(o[g * 8 + 7]) * inv_l);
          *reinterpret_cast<uint4*>(orow + c * 32 + g * 8) = u;
        }
      }
    }
  }

  vc::tc_fence_before();
  __syncthreads();
  if (warp == 2) vc::tmem_dealloc(tmem_base, 512);
}

template <bool kTuned, int kPoly>
int launch_sa(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AttnParams& p, dim3 grid,
              cudaStream_t s) {
  VC_ENSURE_DYN_SMEM((flash_attn_sa_kernel<kTuned, kPoly>), SA_SMEM_BYTES, "vc_flash_attn_fwd");
  flash_attn_sa_kernel<kTuned, kPoly><<<grid, NUM_THREADS, SA_SMEM_BYTES, s>>>(tq, tk, tv, p);
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// Variant 6 ("x1", short key sequences = cross-attention to the 512 text tokens, K7).
// With Lk = 512 a 256-row CTA of the kernels above has only 4 K/V tiles of work (~8k tensor cycles) between a prologue
// (barrier init, TMEM alloc, 64 KB of Q from HBM) and an epilogue (64 KB of O) that nothing overlaps: one CTA per SM
// (224 KB smem, 512 TMEM columns) ran cross-attention at 35 % of the tensor peak.  Here a CTA is half the size —
// ONE 128-row Q tile, S (128 cols) | O (128 cols) = 256 TMEM columns, Q + a 2-slot K/V ring = 96 KB smem, 192 threads —
// so TWO CTAs are resident per SM and one's prologue / softmax / epilogue hides behind the other's MMAs.
//   warp 0  TMA producer (Q, then K_0 V_0 K_1 V_1 ... alternating over the two slots)
//   warp 1  TMEM allocator, then one thread issues S_j = Q K_j^T and O += P_j V_j (P read from TMEM over S, TS MMA)
//   warps 2-5  softmax, thread = query row = TMEM lane (lane quarter = warp & 3)
// No setmaxnreg here: it is a warpgroup-wide instruction and warps 0-3 of this 192-thread CTA have different roles;
// __launch_bounds__(192, 2) already gives every thread 168 registers.
// ---------------------------------------------------------------------------------------------------------------
constexpr int X1_THREADS = 192;
constexpr int X1_NSLOT = 2;
constexpr int X1_SMEM_BYTES = TILE_BYTES + X1_NSLOT * TILE_BYTES + 256 + 1024;

__global__ void __launch_bounds__(X1_THREADS, 2)
flash_attn_x1_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                     const __grid_constant__ CUtensorMap tmV, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sKV = smem + TILE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + X1_NSLOT * TILE_BYTES);
  uint64_t* q_full = bars;                  // 1
  uint64_t* kv_full = bars + 1;             // 2
  uint64_t* kv_empty = kv_full + X1_NSLOT;  // 2
  uint64_t* s_full = kv_empty + X1_NSLOT;   // 1
  uint64_t* p_ready = s_full + 1;           // 1
  uint64_t* o_full = p_ready + 1;           // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int q_blk = blockIdx.x, head = blockIdx.y, batch = blockIdx.z;
  const int nk = (p.Lk + BKV - 1) / BKV;

  if (threadIdx.x == 0) {
    vc::mbar_init(q_full, 1);
    for (int i = 0; i < X1_NSLOT; ++i) {
      vc::mbar_init(&kv_full[i], 1);
      vc::mbar_init(&kv_empty[i], 1);
    }
    vc::mbar_init(s_full, 1);
    vc::mbar_init(p_ready, 128);
    vc::mbar_init(o_full, 1);
    vc::fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    vc::tma_prefetch_desc(&tmQ);
    vc::tma_prefetch_desc(&tmK);
    vc::tma_prefetch_desc(&tmV);
  }
  if (warp == 1) vc::tmem_alloc(tmem_slot, 256);
  vc::tc_fence_before();
  __syncthreads();
  vc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 2) {
    if (warp == 0) {
      if (vc::elect_one()) {
        vc::mbar_expect_tx(q_full, TILE_BYTES);
#pragma unroll
        for (int h = 0; h < 2; ++h) vc::tma_load_4d(sQ + h * HALF_BYTES, &tmQ, q_full, h * 64, q_blk * BQ, head, batch);
        for (int idx = 0; idx < 2 * nk; ++idx) {
          const int slot = idx & 1;
          const uint32_t ph = (idx >> 1) & 1;
          vc::mbar_wait(&kv_empty[slot], ph ^ 1);
          vc::mbar_expect_tx(&kv_full[slot], TILE_BYTES);
          const CUtensorMap* tm = (idx & 1) ? &tmV : &tmK;
          const int row0 = (idx >> 1) * BKV;
#pragma unroll
          for (int h = 0; h < 2; ++h)
            vc::tma_load_4d(sKV + slot * TILE_BYTES + h * HALF_BYTES, tm, &kv_full[slot], h * 64, row0, head, batch);
        }
      }
    } else {
      if (vc::elect_one()) {
        constexpr uint32_t idesc_s = vc::umma_idesc_bf16(BQ, BKV, 0, 0);  // Q (K-major) x K (K-major)
        constexpr uint32_t idesc_o = vc::umma_idesc_bf16(BQ, HD, 0, 1);   // P (TMEM) x V (MN-major)
        const uint32_t qa = vc::smem_u32(sQ);
        const uint32_t ka = vc::smem_u32(sKV), va = vc::smem_u32(sKV + TILE_BYTES);   // K tiles: slot 0, V tiles: slot 1
        auto issue_s = [&]() {
#pragma unroll
          for (int k = 0; k < HD / 16; ++k) {
            const uint32_t off = (k >> 2) * HALF_BYTES + (k & 3) * 32;
            vc::umma_ss(tmem_base, vc::umma_desc_sw128(qa + off, 16, 1024), vc::umma_desc_sw128(ka + off, 16, 1024),
                        idesc_s, k != 0 ? 1u : 0u);
          }
          vc::umma_commit(s_full);
          vc::umma_commit(&kv_empty[0]);
        };
        vc::mbar_wait(q_full, 0);
        vc::mbar_wait(&kv_full[0], 0);
        vc::tc_fence_after();
        issue_s();
        for (int j = 0; j < nk; ++j) {
          vc::mbar_wait(&kv_full[1], j & 1);   // V_j
          vc::mbar_wait(p_ready, j & 1);       // P_j written over S_j
          vc::tc_fence_after();
#pragma unroll
          for (int k = 0; k < BKV / 16; ++k)
            vc::umma_ts(tmem_base + 128, tmem_base + k * 8, vc::umma_desc_sw128(va + k * 2048, HALF_BYTES, 1024), idesc_o,
                        (j != 0 || k != 0) ? 1u : 0u);
          vc::umma_commit(&kv_empty[1]);
          if (j + 1 < nk) {
            vc::mbar_wait(&kv_full[0], (j + 1) & 1);   // K_{j+1}
            vc::tc_fence_after();
            issue_s();                                 // in-order after PV_j: P_j is consumed before S is overwritten
          } else {
            vc::umma_commit(o_full);
          }
        }
      }
    }
  } else {
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t lane_bits = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t s_addr = tmem_base + lane_bits;
    const uint32_t o_addr = tmem_base + 128 + lane_bits;
    const float sl2 = p.scale_log2;
    float m_run = -INFINITY, l_run = 0.f;

    for (int j = 0; j < nk; ++j) {
      vc::mbar_wait(s_full, j & 1);
      vc::tc_fence_after();
      uint32_t s[128];
#pragma unroll
      for (int c = 0; c < 4; ++c) vc::tmem_ld32(s_addr + c * 32, s + c * 32);
      vc::tmem_ld_wait();
      if (j == nk - 1) {
        const int valid = p.Lk - j * BKV;
        if (valid < BKV) {
#pragma unroll
          for (int i = 0; i < 128; ++i)
            if (i >= valid) s[i] = __float_as_uint(-INFINITY);
        }
      }
      float mx[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) mx[i] = __uint_as_float(s[i]);
#pragma unroll
      for (int i = 8; i < 128; ++i) mx[i & 7] = fmaxf(mx[i & 7], __uint_as_float(s[i]));
      const float mt = fmaxf(fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])), fmaxf(fmaxf(mx[4], mx[5]), fmaxf(mx[6], mx[7])));
      const float m_new = fmaxf(m_run, mt);
      const bool need = (m_new - m_run) * sl2 > RESCALE_THRESHOLD;
      if (__any_sync(0xffffffffu, need)) {
        const float alpha = fast_exp2((m_run - m_new) * sl2);
        l_run *= alpha;
        m_run = m_new;
        if (j > 0) {   // S_j done implies PV_{j-1} done; PV_j waits for our p_ready arrival: O is idle
#pragma unroll 1
          for (int c = 0; c < 4; ++c) {
            uint32_t o[32];
            vc::tmem_ld32(o_addr + c * 32, o);
            vc::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            vc::tmem_st32(o_addr + c * 32, o);
          }
          vc::tmem_st_wait();
        }
      }
      const float mb = m_run * sl2;
      float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
      for (int i = 0; i < 64; ++i) {
        const float p0 = fast_exp2(fmaf(__uint_as_float(s[2 * i]), sl2, -mb));
        const float p1 = fast_exp2(fmaf(__uint_as_float(s[2 * i + 1]), sl2, -mb));
        sum0 += p0;
        sum1 += p1;
        s[i] = vc::pack_bf16x2(p0, p1);
      }
      vc::tmem_st32(s_addr, s);
      vc::tmem_st32(s_addr + 32, s + 32);
      vc::tmem_st_wait();
      l_run += sum0 + sum1;
      vc::tc_fence_before();
      vc::mbar_arrive(p_ready);
    }

    vc::mbar_wait(o_full, 0);
    vc::tc_fence_after();
    const float inv_l = 1.0f / l_run;
    const int q_row = q_blk * BQ + row;
    __nv_bfloat16* orow = p.O + (long long)batch * p.o_stride_b + (long long)q_row * p.o_stride_l +
                          (long long)head * p.o_stride_h;
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      uint32_t o[32];
      vc::tmem_ld32(o_addr + c * 32, o);
      vc::tmem_ld_wait();
      if (q_row < p.Lq) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 u;
          u.x = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 0]) * inv_l, __uint_as_float(o[g * 8 + 1]) * inv_l);
          u.y = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 2]) * inv_l, __uint_as_float(o[g * 8 + 3]) * inv_l);
          u.z = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 4]) * inv_l, __uint_as_float(o[g * 8 + 5]) * inv_l);
          u.w = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 6]) * inv_l, __uint_as_float(o[g * 8 + 7]) * inv_l);
          *reinterpret_cast<uint4*>(orow + c * 32 + g * 8) = u;
        }
      }
    }
  }

  vc::tc_fence_before();
  __syncthreads();
  if (warp == 1) vc::tmem_dealloc(tmem_base, 256);
}

int launch_x1(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AttnParams& p, int B, int H,
              cudaStream_t s) {
  VC_ENSURE_DYN_SMEM(flash_attn_x1_kernel, X1_SMEM_BYTES, "vc_flash_attn_fwd");
  {
    static bool carve[vc_host::kMaxDevices] = {};
    const int dev = vc_host::current_device();
    if (!carve[dev]) {  // two CTAs of 96 KB per SM need the full shared-memory carve-out
      cudaError_t e = cudaFuncSetAttribute(flash_attn_x1_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
      if (e != cudaSuccess) {
        vc_host::set_error("vc_flash_attn_fwd: cannot set the shared-memory carve-out: %s", cudaGetErrorString(e));
        return (int)e;
      }
      carve[dev] = true;
    }
  }
  dim3 grid((p.Lq + BQ - 1) / BQ, H, B);
  flash_attn_x1_kernel<<<grid, X1_THREADS, X1_SMEM_BYTES, s>>>(tq, tk, tv, p);
  return 0;
}


// ---------------------------------------------------------------------------------------------------------------
// Variant 7 ("x2", short key sequences, round 2).  Variant 6 keeps S -> softmax -> PV of one key tile strictly in series
// (P aliases S, so S_{j+1} may only be issued after PV_j): with four 128-key tiles a CTA is a latency chain of
// 4 x (512 + ~2500 + 512) cycles and two CTAs per SM only halve it (ncu, profiles/ncu_xattn_x1_r01j_summary.txt: tensor
// pipe 33 % active, issue-active 27 %).  Here the key tile is 64 wide and S is double-buffered inside the same 256 TMEM
// columns:   S0 (64) | S1 (64) | O (128),   P_j written over the first 32 columns of S_{j&1}
// MMA issue order (one thread, in-order pipe):   S_0 S_1 | PV_0 S_2 | PV_1 S_3 | PV_2 S_4 | ...
// so S_{j+1} is complete before softmax_j ends and the softmax warps never wait for the tensor pipe; the MUFU (exp2) and
// the tensor pipe overlap instead of alternating, and the per-thread working set halves (64 scores).  Same CTA shape as
// variant 6 (one 128-row Q tile, 192 threads, 96 KB smem: Q + two 16 KB K slots + two 16 KB V slots), two CTAs per SM.
// A lazy O rescale at tile j needs O idle: PV_{j-1} may still be running when S_j completes (it was issued after S_j),
// so the softmax warps wait for its commit (pv_done) first; PV_j cannot start before they arrive on p_ready[j&1].
// ---------------------------------------------------------------------------------------------------------------
constexpr int X2_BKV = 64;
constexpr int X2_KV_BYTES = X2_BKV * HD * 2;        // 16 KB: two 64-column halves of 64 rows x 128 B
constexpr int X2_KV_HALF = X2_BKV * 128;            // 8 KB
constexpr int X2_SMEM_BYTES = TILE_BYTES + 4 * X2_KV_BYTES + 256 + 1024;

template <int kPoly>
__global__ void __launch_bounds__(X1_THREADS, 2)
flash_attn_x2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                     const __grid_constant__ CUtensorMap tmV, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = smem + TILE_BYTES;               // 2 slots
  uint8_t* sV = sK + 2 * X2_KV_BYTES;            // 2 slots
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + 2 * X2_KV_BYTES);
  uint64_t* q_full = bars;            // 1
  uint64_t* k_full = bars + 1;        // 2
  uint64_t* k_empty = k_full + 2;     // 2
  uint64_t* v_full = k_empty + 2;     // 2
  uint64_t* v_empty = v_full + 2;     // 2
  uint64_t* s_full = v_empty + 2;     // 2
  uint64_t* p_ready = s_full + 2;     // 2
  uint64_t* pv_done = p_ready + 2;    // 2
  uint64_t* o_full = pv_done + 2;     // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int q_blk = blockIdx.x, head = blockIdx.y, batch = blockIdx.z;
  const int nk = (p.Lk + X2_BKV - 1) / X2_BKV;

  if (threadIdx.x == 0) {
    vc::mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      vc::mbar_init(&k_full[i], 1);
      vc::mbar_init(&k_empty[i], 1);
      vc::mbar_init(&v_full[i], 1);
      vc::mbar_init(&v_empty[i], 1);
      vc::mbar_init(&s_full[i], 1);
      vc::mbar_init(&p_ready[i], 128);
      vc::mbar_init(&pv_done[i], 1);
    }
    vc::mbar_init(o_full, 1);
    vc::fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    vc::tma_prefetch_desc(&tmQ);
    vc::tma_prefetch_desc(&tmK);
    vc::tma_prefetch_desc(&tmV);
  }
  if (warp == 1) vc::tmem_alloc(tmem_slot, 256);
  vc::tc_fence_before();
  __syncthreads();
  vc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 2) {
    if (warp == 0) {
      if (vc::elect_one()) {
        vc::mbar_expect_tx(q_full, TILE_BYTES);
#pragma unroll
        for (int h = 0; h < 2; ++h) vc::tma_load_4d(sQ + h * HALF_BYTES, &tmQ, q_full, h * 64, q_blk * BQ, head, batch);
        // first-use order: K_0 K_1 V_0 K_2 V_1 K_3 V_2 ...
        auto load_k = [&](int j) {
          const int slot = j & 1;
          vc::mbar_wait(&k_empty[slot], ((j >> 1) & 1) ^ 1);
          vc::mbar_expect_tx(&k_full[slot], X2_KV_BYTES);
#pragma unroll
          for (int h = 0; h < 2; ++h)
            vc::tma_load_4d(sK + slot * X2_KV_BYTES + h * X2_KV_HALF, &tmK, &k_full[slot], h * 64, j * X2_BKV, head, batch);
        };
        auto load_v = [&](int j) {
          const int slot = j & 1;
          vc::mbar_wait(&v_empty[slot], ((j >> 1) & 1) ^ 1);
          vc::mbar_expect_tx(&v_full[slot], X2_KV_BYTES);
#pragma unroll
          for (int h = 0; h < 2; ++h)
            vc::tma_load_4d(sV + slot * X2_KV_BYTES + h * X2_KV_HALF, &tmV, &v_full[slot], h * 64, j * X2_BKV, head, batch);
        };
        load_k(0);
        if (nk > 1) load_k(1);
        for (int j = 0; j < nk; ++j) {
          load_v(j);
          if (j + 2 < nk) load_k(j + 2);
        }
      }
    } else {
      if (vc::elect_one()) {
        constexpr uint32_t idesc_s = vc::umma_idesc_bf16(BQ, X2_BKV, 0, 0);  // Q (K-major) x K_j (K-major): 128 x 64
        constexpr uint32_t idesc_o = vc::umma_idesc_bf16(BQ, HD, 0, 1);      // P (TMEM) x V_j (MN-major): 128 x 128, K = 64
        const uint32_t qa = vc::smem_u32(sQ);
        auto issue_s = [&](int j) {
          const int b = j & 1;
          vc::mbar_wait(&k_full[b], (j >> 1) & 1);
          vc::tc_fence_after();
          const uint32_t ka = vc::smem_u32(sK + b * X2_KV_BYTES);
#pragma unroll
          for (int k = 0; k < HD / 16; ++k) {
            const uint32_t qoff = (k >> 2) * HALF_BYTES + (k & 3) * 32;
            const uint32_t koff = (k >> 2) * X2_KV_HALF + (k & 3) * 32;
            vc::umma_ss(tmem_base + b * X2_BKV, vc::umma_desc_sw128(qa + qoff, 16, 1024),
                        vc::umma_desc_sw128(ka + koff, 16, 1024), idesc_s, k != 0 ? 1u : 0u);
          }
          vc::umma_commit(&s_full[b]);
          vc::umma_commit(&k_empty[b]);
        };
        vc::mbar_wait(q_full, 0);
        issue_s(0);
        if (nk > 1) issue_s(1);
        for (int j = 0; j < nk; ++j) {
          const int b = j & 1;
          vc::mbar_wait(&v_full[b], (j >> 1) & 1);    // V_j
          vc::mbar_wait(&p_ready[b], (j >> 1) & 1);   // P_j written over S_j
          vc::tc_fence_after();
          const uint32_t va = vc::smem_u32(sV + b * X2_KV_BYTES);
#pragma unroll
          for (int k = 0; k < X2_BKV / 16; ++k)
            vc::umma_ts(tmem_base + 128, tmem_base + b * X2_BKV + k * 8, vc::umma_desc_sw128(va + k * 2048, X2_KV_HALF, 1024),
                        idesc_o, (j != 0 || k != 0) ? 1u : 0u);
          vc::umma_commit(&v_empty[b]);
          vc::umma_commit(&pv_done[b]);
          if (j + 2 < nk) issue_s(j + 2);   // in-order after PV_j: P_j is consumed before S_{j+2} overwrites its buffer
        }
        vc::umma_commit(o_full);
      }
    }
  } else {
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t lane_bits = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t o_addr = tmem_base + 128 + lane_bits;
    const float sl2 = p.scale_log2;
    float m_run = -INFINITY, l_run = 0.f;

    for (int j = 0; j < nk; ++j) {
      const int b = j & 1;
      const uint32_t s_addr = tmem_base + b * X2_BKV + lane_bits;
      vc::mbar_wait(&s_full[b], (j >> 1) & 1);
      vc::tc_fence_after();
      uint32_t s[64];
      vc::tmem_ld32(s_addr, s);
      vc::tmem_ld32(s_addr + 32, s + 32);
      vc::tmem_ld_wait();
      if (j == nk - 1) {
        const int valid = p.Lk - j * X2_BKV;
        if (valid < X2_BKV) {
#pragma unroll
          for (int i = 0; i < 64; ++i)
            if (i >= valid) s[i] = __float_as_uint(-INFINITY);
        }
      }
      float mx[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) mx[i] = __uint_as_float(s[i]);
#pragma unroll
      for (int i = 8; i < 64; ++i) mx[i & 7] = fmaxf(mx[i & 7], __uint_as_float(s[i]));
      const float mt = fmaxf(fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])), fmaxf(fmaxf(mx[4], mx[5]), fmaxf(mx[6], mx[7])));
      const float m_new = fmaxf(m_run, mt);
      const bool need = (m_new - m_run) * sl2 > RESCALE_THRESHOLD;
      if (__any_sync(0xffffffffu, need)) {
        const float alpha = fast_exp2((m_run - m_new) * sl2);
        l_run *= alpha;
        m_run = m_new;
        if (j > 0) {
          // O must be idle: PV_{j-1} was issued after S_j and may still run; PV_j waits for our p_ready arrival
          vc::mbar_wait(&pv_done[(j - 1) & 1], ((j - 1) >> 1) & 1);
          vc::tc_fence_after();
#pragma unroll 1
          for (int c = 0; c < 4; ++c) {
            uint32_t o[32];
            vc::tmem_ld32(o_addr + c * 32, o);
            vc::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            vc::tmem_st32(o_addr + c * 32, o);
          }
          vc::tmem_st_wait();
        }
      }
      const float mb = m_run * sl2;
      float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        float p0, p1;
        if (i < kPoly / 2) {   // a share of the exponentials on the FMA pipe (the MUFU is the shared bottleneck of 2 CTAs)
          exp2_poly2(p0, p1, fmaf(__uint_as_float(s[2 * i]), sl2, -mb), fmaf(__uint_as_float(s[2 * i + 1]), sl2, -mb));
        } else {
          p0 = fast_exp2(fmaf(__uint_as_float(s[2 * i]), sl2, -mb));
          p1 = fast_exp2(fmaf(__uint_as_float(s[2 * i + 1]), sl2, -mb));
        }
        sum0 += p0;
        sum1 += p1;
        s[i] = vc::pack_bf16x2(p0, p1);
      }
      vc::tmem_st32(s_addr, s);
      vc::tmem_st_wait();
      l_run += sum0 + sum1;
      vc::tc_fence_before();
      vc::mbar_arrive(&p_ready[b]);
    }

    vc::mbar_wait(o_full, 0);
    vc::tc_fence_after();
    const float inv_l = 1.0f / l_run;
    const int q_row = q_blk * BQ + row;
    __nv_bfloat16* orow = p.O + (long long)batch * p.o_stride_b + (long long)q_row * p.o_stride_l +
                          (long long)head * p.o_stride_h;
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      uint32_t o[32];
      vc::tmem_ld32(o_addr + c * 32, o);
      vc::tmem_ld_wait();
      if (q_row < p.Lq) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 u;
          u.x = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 0]) * inv_l, __uint_as_float(o[g * 8 + 1]) * inv_l);
          u.y = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 2]) * inv_l, __uint_as_float(o[g * 8 + 3]) * inv_l);
          u.z = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 4]) * inv_l, __uint_as_float(o[g * 8 + 5]) * inv_l);
          u.w = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 6]) * inv_l, __uint_as_float(o[g * 8 + 7]) * inv_l);
          *reinterpret_cast<uint4*>(orow + c * 32 + g * 8) = u;
        }
      }
    }
  }

  vc::tc_fence_before();
  __syncthreads();
  if (warp == 1) vc::tmem_dealloc(tmem_base, 256);
}

template <int kPoly>
int launch_x2(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AttnParams& p, int B, int H,
              cudaStream_t s) {
  VC_ENSURE_DYN_SMEM((flash_attn_x2_kernel<kPoly>), X2_SMEM_BYTES, "vc_flash_attn_fwd");
  {
    static bool carve[vc_host::kMaxDevices] = {};
    const int dev = vc_host::current_device();
    if (!carve[dev]) {  // two CTAs of 96 KB per SM need the full shared-memory carve-out
      cudaError_t e = cudaFuncSetAttribute(flash_attn_x2_kernel<kPoly>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
      if (e != cudaSuccess) {
        vc_host::set_error("vc_flash_attn_fwd: cannot set the shared-memory carve-out: %s", cudaGetErrorString(e));
        return (int)e;
      }
      carve[dev] = true;
    }
  }
  dim3 grid((p.Lq + BQ - 1) / BQ, H, B);
  flash_attn_x2_kernel<kPoly><<<grid, X1_THREADS, X2_SMEM_BYTES, s>>>(tq, tk, tv, p);
  return 0;
}


// ---------------------------------------------------------------------------------------------------------------
// Variant 9 ("x3", persistent short-key kernel, round 2).  Timing variant 7 at Lk = 192 vs 512 showed 1.3 k cycles per
// 64-key step but 12.6 k cycles of un-overlapped prologue + epilogue per CTA (barrier init, TMEM allocation, Q from HBM,
// O to HBM, launch): 54 % of the kernel at Lk = 512.  Here the same pipeline (64-key tiles, S0 | S1 | O in 256 TMEM
// columns, two CTAs per SM) runs in a PERSISTENT CTA that walks Q tiles t = blockIdx.x, blockIdx.x + gridDim.x, ...:
//   * barriers and TMEM are set up once;
//   * the single Q buffer is refilled for tile t+1 as soon as the last S MMA of tile t has read it (q_empty = commit after
//     S_{n-1}), i.e. about two softmax steps + the epilogue ahead of its first use;
//   * the MMA warp issues S_0 / S_1 of tile t+1 while the softmax warps still drain O of tile t; only PV_0 of tile t+1
//     (which overwrites O) waits for that epilogue (o_empty).
// All ring indices run on global counters (g = t_local * nk + j), so odd tile counts need no special case.
// ---------------------------------------------------------------------------------------------------------------
template <int kPoly>
__global__ void __launch_bounds__(X1_THREADS, 2)
flash_attn_x3_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                     const __grid_constant__ CUtensorMap tmV, const AttnParams p, int H, int n_qblk, int n_tiles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = smem + TILE_BYTES;               // 2 slots
  uint8_t* sV = sK + 2 * X2_KV_BYTES;            // 2 slots
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + 2 * X2_KV_BYTES);
  uint64_t* q_full = bars;            // 1
  uint64_t* q_empty = bars + 1;       // 1
  uint64_t* k_full = bars + 2;        // 2
  uint64_t* k_empty = k_full + 2;     // 2
  uint64_t* v_full = k_empty + 2;     // 2
  uint64_t* v_empty = v_full + 2;     // 2
  uint64_t* s_full = v_empty + 2;     // 2
  uint64_t* p_ready = s_full + 2;     // 2
  uint64_t* pv_done = p_ready + 2;    // 2
  uint64_t* o_full = pv_done + 2;     // 1
  uint64_t* o_empty = o_full + 1;     // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_empty + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int nk = (p.Lk + X2_BKV - 1) / X2_BKV;

  if (threadIdx.x == 0) {
    vc::mbar_init(q_full, 1);
    vc::mbar_init(q_empty, 1);
    for (int i = 0; i < 2; ++i) {
      vc::mbar_init(&k_full[i], 1);
      vc::mbar_init(&k_empty[i], 1);
      vc::mbar_init(&v_full[i], 1);
      vc::mbar_init(&v_empty[i], 1);
      vc::mbar_init(&s_full[i], 1);
      vc::mbar_init(&p_ready[i], 128);
      vc::mbar_init(&pv_done[i], 1);
    }
    vc::mbar_init(o_full, 1);
    vc::mbar_init(o_empty, 128);
    vc::fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    vc::tma_prefetch_desc(&tmQ);
    vc::tma_prefetch_desc(&tmK);
    vc::tma_prefetch_desc(&tmV);
  }
  if (warp == 1) vc::tmem_alloc(tmem_slot, 256);
  vc::tc_fence_before();
  __syncthreads();
  vc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // tile -> (query block, head, batch): query blocks of one (batch, head) are consecutive so co-running CTAs share K/V in L2
  auto decode = [&](int t, int& q_blk, int& head, int& batch) {
    q_blk = t % n_qblk;
    const int bh = t / n_qblk;
    head = bh % H;
    batch = bh / H;
  };

  if (warp == 0) {
    if (vc::elect_one()) {
      int tl = 0;  // local tile counter
      for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++tl) {
        int q_blk, head, batch;
        decode(t, q_blk, head, batch);
        const int g0 = tl * nk;
        auto load_k = [&](int j) {
          const int g = g0 + j, slot = g & 1;
          vc::mbar_wait(&k_empty[slot], ((g >> 1) & 1) ^ 1);
          vc::mbar_expect_tx(&k_full[slot], X2_KV_BYTES);
#pragma unroll
          for (int h = 0; h < 2; ++h)
            vc::tma_load_4d(sK + slot * X2_KV_BYTES + h * X2_KV_HALF, &tmK, &k_full[slot], h * 64, j * X2_BKV, head, batch);
        };
        auto load_v = [&](int j) {
          const int g = g0 + j, slot = g & 1;
          vc::mbar_wait(&v_empty[slot], ((g >> 1) & 1) ^ 1);
          vc::mbar_expect_tx(&v_full[slot], X2_KV_BYTES);
#pragma unroll
          for (int h = 0; h < 2; ++h)
            vc::tma_load_4d(sV + slot * X2_KV_BYTES + h * X2_KV_HALF, &tmV, &v_full[slot], h * 64, j * X2_BKV, head, batch);
        };
        vc::mbar_wait(q_empty, (tl & 1) ^ 1);   // the S MMAs of the previous tile have read Q
        vc::mbar_expect_tx(q_full, TILE_BYTES);
#pragma unroll
        for (int h = 0; h < 2; ++h) vc::tma_load_4d(sQ + h * HALF_BYTES, &tmQ, q_full, h * 64, q_blk * BQ, head, batch);
        load_k(0);
        if (nk > 1) load_k(1);
        for (int j = 0; j < nk; ++j) {
          load_v(j);
          if (j + 2 < nk) load_k(j + 2);
        }
      }
    }
  } else if (warp == 1) {
    if (vc::elect_one()) {
      constexpr uint32_t idesc_s = vc::umma_idesc_bf16(BQ, X2_BKV, 0, 0);
      constexpr uint32_t idesc_o = vc::umma_idesc_bf16(BQ, HD, 0, 1);
      const uint32_t qa = vc::smem_u32(sQ);
      int tl = 0;
      for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++tl) {
        const int g0 = tl * nk;
        auto issue_s = [&](int j) {
          const int g = g0 + j, b = g & 1;
          vc::mbar_wait(&k_full[b], (g >> 1) & 1);
          vc::tc_fence_after();
          const uint32_t ka = vc::smem_u32(sK + b * X2_KV_BYTES);
#pragma unroll
          for (int k = 0; k < HD / 16; ++k) {
            const uint32_t qoff = (k >> 2) * HALF_BYTES + (k & 3) * 32;
            const uint32_t koff = (k >> 2) * X2_KV_HALF + (k & 3) * 32;
            vc::umma_ss(tmem_base + b * X2_BKV, vc::umma_desc_sw128(qa + qoff, 16, 1024),
                        vc::umma_desc_sw128(ka + koff, 16, 1024), idesc_s, k != 0 ? 1u : 0u);
          }
          vc::umma_commit(&s_full[b]);
          vc::umma_commit(&k_empty[b]);
          if (j == nk - 1) vc::umma_commit(q_empty);   // last reader of this tile's Q
        };
        vc::mbar_wait(q_full, tl & 1);
        issue_s(0);
        if (nk > 1) issue_s(1);
        for (int j = 0; j < nk; ++j) {
          const int g = g0 + j, b = g & 1;
          vc::mbar_wait(&v_full[b], (g >> 1) & 1);
          vc::mbar_wait(&p_ready[b], (g >> 1) & 1);
          if (j == 0) vc::mbar_wait(o_empty, (tl & 1) ^ 1);   // the previous tile's O has been read out
          vc::tc_fence_after();
          const uint32_t va = vc::smem_u32(sV + b * X2_KV_BYTES);
#pragma unroll
          for (int k = 0; k < X2_BKV / 16; ++k)
            vc::umma_ts(tmem_base + 128, tmem_base + b * X2_BKV + k * 8, vc::umma_desc_sw128(va + k * 2048, X2_KV_HALF, 1024),
                        idesc_o, (j != 0 || k != 0) ? 1u : 0u);
          vc::umma_commit(&v_empty[b]);
          vc::umma_commit(&pv_done[b]);
          if (j + 2 < nk) issue_s(j + 2);
        }
        vc::umma_commit(o_full);
      }
    }
  } else {
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t lane_bits = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t o_addr = tmem_base + 128 + lane_bits;
    const float sl2 = p.scale_log2;
    int tl = 0;
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++tl) {
      int q_blk, head, batch;
      decode(t, q_blk, head, batch);
      const int g0 = tl * nk;
      float m_run = -INFINITY, l_run = 0.f;
      for (int j = 0; j < nk; ++j) {
        const int g = g0 + j, b = g & 1;
        const uint32_t s_addr = tmem_base + b * X2_BKV + lane_bits;
        vc::mbar_wait(&s_full[b], (g >> 1) & 1);
        vc::tc_fence_after();
        uint32_t s[64];
        vc::tmem_ld32(s_addr, s);
        vc::tmem_ld32(s_addr + 32, s + 32);
        vc::tmem_ld_wait();
        if (j == nk - 1) {
          const int valid = p.Lk - j * X2_BKV;
          if (valid < X2_BKV) {
#pragma unroll
            for (int i = 0; i < 64; ++i)
              if (i >= valid) s[i] = __float_as_uint(-INFINITY);
          }
        }
        float mx[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) mx[i] = __uint_as_float(s[i]);
#pragma unroll
        for (int i = 8; i < 64; ++i) mx[i & 7] = fmaxf(mx[i & 7], __uint_as_float(s[i]));
        const float mt = fmaxf(fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])), fmaxf(fmaxf(mx[4], mx[5]), fmaxf(mx[6], mx[7])));
        const float m_new = fmaxf(m_run, mt);
        const bool need = (m_new - m_run) * sl2 > RESCALE_THRESHOLD;
        if (__any_sync(0xffffffffu, need)) {
          const float alpha = fast_exp2((m_run - m_new) * sl2);
          l_run *= alpha;
          m_run = m_new;
          if (j > 0) {
            const int gp = g - 1;   // O must be idle: PV_{j-1} may still run; PV_j waits for our p_ready arrival
            vc::mbar_wait(&pv_done[gp & 1], (gp >> 1) & 1);
            vc::tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
              uint32_t o[32];
              vc::tmem_ld32(o_addr + c * 32, o);
              vc::tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
              vc::tmem_st32(o_addr + c * 32, o);
            }
            vc::tmem_st_wait();
          }
        }
        const float mb = m_run * sl2;
        float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float p0, p1;
          if (i < kPoly / 2) {
            exp2_poly2(p0, p1, fmaf(__uint_as_float(s[2 * i]), sl2, -mb), fmaf(__uint_as_float(s[2 * i + 1]), sl2, -mb));
          } else {
            p0 = fast_exp2(fmaf(__uint_as_float(s[2 * i]), sl2, -mb));
            p1 = fast_exp2(fmaf(__uint_as_float(s[2 * i + 1]), sl2, -mb));
          }
          sum0 += p0;
          sum1 += p1;
          s[i] = vc::pack_bf16x2(p0, p1);
        }
        vc::tmem_st32(s_addr, s);
        vc::tmem_st_wait();
        l_run += sum0 + sum1;
        vc::tc_fence_before();
        vc::mbar_arrive(&p_ready[b]);
      }

      vc::mbar_wait(o_full, tl & 1);
      vc::tc_fence_after();
      const float inv_l = 1.0f / l_run;
      const int q_row = q_blk * BQ + row;
      __nv_bfloat16* orow = p.O + (long long)batch * p.o_stride_b + (long long)q_row * p.o_stride_l +
                            (long long)head * p.o_stride_h;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t o[32];
        vc::tmem_ld32(o_addr + c * 32, o);
        vc::tmem_ld_wait();
        if (c == 3) {   // O is in registers: the next tile's PV_0 may overwrite the accumulator
          vc::tc_fence_before();
          vc::mbar_arrive(o_empty);
        }
        if (q_row < p.Lq) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            uint4 u;
            u.x = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 0]) * inv_l, __uint_as_float(o[g * 8 + 1]) * inv_l);
            u.y = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 2]) * inv_l, __uint_as_float(o[g * 8 + 3]) * inv_l);
            u.z = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 4]) * inv_l, __uint_as_float(o[g * 8 + 5]) * inv_l);
            u.w = vc::pack_bf16x2(__uint_as_float(o[g * 8 + 6]) * inv_l, __uint_as_float(o[g * 8 + 7]) * inv_l);
            *reinterpret_cast<uint4*>(orow + c * 32 + g * 8) = u;
          }
        }
      }
    }
  }

  vc::tc_fence_before();
  __syncthreads();
  if (warp == 1) vc::tmem_dealloc(tmem_base, 256);
}

template <int kPoly>
int launch_x3(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AttnParams& p, int B, int H,
              cudaStream_t s) {
  VC_ENSURE_DYN_SMEM((flash_attn_x3_kernel<kPoly>), X2_SMEM_BYTES, "vc_flash_attn_fwd");
  {
    static bool carve[vc_host::kMaxDevices] = {};
    const int dev = vc_host::current_device();
    if (!carve[dev]) {
      cudaError_t e = cudaFuncSetAttribute(flash_attn_x3_kernel<kPoly>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
      if (e != cudaSuccess) {
        vc_host::set_error("vc_flash_attn_fwd: cannot set the shared-memory carve-out: %s", cudaGetErrorString(e));
        return (int)e;
      }
      carve[dev] = true;
    }
  }
  const int n_qblk = (p.Lq + BQ - 1) / BQ;
  const long long tiles = (long long)n_qblk * H * B;
  const int cap = 2 * vc_host::sm_count();   // two resident CTAs per SM
  const int grid = (int)(tiles < cap ? tiles : cap);
  flash_attn_x3_kernel<kPoly><<<grid, X1_THREADS, X2_SMEM_BYTES, s>>>(tq, tk, tv, p, H, n_qblk, (int)tiles);
  return 0;
}

}  // namespace

// strides are in ELEMENTS: (batch, token, head); the head_dim axis must be contiguous (stride 1), head_dim == 128.
static int attn_impl(const void* q, const void* k, const void* v, void* o, int B, int H, int Lq, int Lk, int head_dim,
                     const long long* q_strides, const long long* k_strides, const long long* v_strides,
                     const long long* o_strides, float softmax_scale, int variant, void* const* peer_o, int n_peers,
                     int rows_per_peer, void* stream) {
  VC_CHECK_ARG(q && k && v && o && q_strides && k_strides && v_strides && o_strides, "vc_flash_attn_fwd: null pointer");
  VC_CHECK_ARG(head_dim == HD, "vc_flash_attn_fwd: head_dim %d unsupported (only 128)", head_dim);
  VC_CHECK_ARG(B > 0 && H > 0 && Lq >= 0 && Lk > 0, "vc_flash_attn_fwd: bad sizes B=%d H=%d Lq=%d Lk=%d", B, H, Lq, Lk);
  VC_CHECK_ARG(variant >= 0 && variant <= 9,
               "vc_flash_attn_fwd: variant must be 0 (P in smem), 1 (P in TMEM over S), 2-5 (shared S buffer), 6 (short-key x1), "
               "7/8 (short-key x2, double-buffered 64-key S) or 9 (persistent x3)");
  const long long* all[4] = {q_strides, k_strides, v_strides, o_strides};
  for (int t = 0; t < 4; ++t)
    for (int i = 0; i < 3; ++i)
      VC_CHECK_ARG(all[t][i] % 8 == 0 && all[t][i] > 0, "vc_flash_attn_fwd: strides must be positive multiples of 8 elements");
  VC_CHECK_ARG(((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)o) % 16 == 0, "vc_flash_attn_fwd: 16-byte alignment");
  if (Lq == 0) return 0;
  CUtensorMap tq, tk, tv;
  int rc;
  if ((rc = make_qkv_map(&tq, q, B, H, Lq, q_strides[0], q_strides[1], q_strides[2]))) return rc;
  const uint32_t kv_rows = variant >= 7 ? X2_BKV : 128;   // variants 7/8 load 64-key tiles
  if ((rc = make_qkv_map(&tk, k, B, H, Lk, k_strides[0], k_strides[1], k_strides[2], kv_rows))) return rc;
  if ((rc = make_qkv_map(&tv, v, B, H, Lk, v_strides[0], v_strides[1], v_strides[2], kv_rows))) return rc;
  AttnParams p;
  p.O = (__nv_bfloat16*)o;
  p.o_stride_b = o_strides[0]; p.o_stride_l = o_strides[1]; p.o_stride_h = o_strides[2];
  p.Lq = Lq; p.Lk = Lk;
  p.scale_log2 = softmax_scale * 1.4426950408889634f;
  p.n_peers = n_peers;
  p.rows_per_peer = rows_per_peer > 0 ? rows_per_peer : 1;
  for (int i = 0; i < 8; ++i) p.o_peer[i] = (__nv_bfloat16*)(i < n_peers ? peer_o[i] : nullptr);
  dim3 grid((Lq + 2 * BQ - 1) / (2 * BQ), H, B);
  cudaStream_t st = (cudaStream_t)stream;
  VC_CHECK_ARG(variant < 6 || n_peers == 0, "vc_flash_attn_fwd_scatter: variants 6-8 have no scatter epilogue");
  switch (variant) {
    case 6: rc = launch_x1(tq, tk, tv, p, B, H, st); break;
    case 7: rc = launch_x2<0>(tq, tk, tv, p, B, H, st); break;
    case 8: rc = launch_x2<16>(tq, tk, tv, p, B, H, st); break;
    case 9: rc = launch_x3<0>(tq, tk, tv, p, B, H, st); break;
    case 0: rc = launch<false>(tq, tk, tv, p, grid, st); break;
    case 1: rc = launch<true>(tq, tk, tv, p, grid, st); break;
    case 2: rc = launch_sa<false, 0>(tq, tk, tv, p, grid, st); break;
    case 3: rc = launch_sa<true, 0>(tq, tk, tv, p, grid, st); break;
    case 4: rc = launch_sa<true, 16>(tq, tk, tv, p, grid, st); break;
    default: rc = launch_sa<true, 32>(tq, tk, tv, p, grid, st); break;
  }
  if (rc) return rc;
  VC_RETURN_LAUNCH_STATUS("vc_flash_attn_fwd");
}

extern "C" int vc_flash_attn_fwd(const void* q, const void* k, const void* v, void* o, int B, int H, int Lq, int Lk,
                                 int head_dim, const long long* q_strides, const long long* k_strides,
                                 const long long* v_strides, const long long* o_strides, float softmax_scale,
                                 int variant, void* stream) {
  return attn_impl(q, k, v, o, B, H, Lq, Lk, head_dim, q_strides, k_strides, v_strides, o_strides, softmax_scale, variant,
                   nullptr, 0, 0, stream);
}

extern "C" int vc_flash_attn_fwd_scatter(const void* q, const void* k, const void* v, int B, int H, int Lq, int Lk,
                                         int head_dim, const long long* q_strides, const long long* k_strides,
                                         const long long* v_strides, void* const* peer_o, int n_peers,
                                         int rows_per_peer, const long long* o_strides, float softmax_scale,
                                         int variant, void* stream) {
  VC_CHECK_ARG(peer_o && n_peers >= 1 && n_peers <= 8, "vc_flash_attn_fwd_scatter: n_peers must be in 1..8");
  VC_CHECK_ARG(variant >= 2, "vc_flash_attn_fwd_scatter: needs a shared-S variant (>= 2)");
  VC_CHECK_ARG(rows_per_peer > 0 && (long long)rows_per_peer * n_peers >= Lq,
               "vc_flash_attn_fwd_scatter: rows_per_peer * n_peers must cover Lq");
  for (int i = 0; i < n_peers; ++i)
    VC_CHECK_ARG(peer_o[i] && (uintptr_t)peer_o[i] % 16 == 0, "vc_flash_attn_fwd_scatter: bad peer pointer %d", i);
  return attn_impl(q, k, v, peer_o[0], B, H, Lq, Lk, head_dim, q_strides, k_strides, v_strides, o_strides, softmax_scale,
                   variant, peer_o, n_peers, rows_per_peer, stream);
}
