// Patch-embedding Conv3d(kernel = stride = (1,2,2)) as a TMA-fed IMPLICIT GEMM on tcgen05:
//   out[n, :] = sum_{c,ph,pw} x[c, f, 2h'+ph, 2w'+pw] * weight[:, c, 0, ph, pw] + bias,   n = (f*H' + h')*W' + w'
// Replaces nn.Conv3d patch_embedding (wan_transformer3d.py:758-759, used wan_transformer3d_versecrafter.py:340-343) and
// geoada_patch_embedding (wan_transformer3d_versecrafter.py:199-201, used :262-267) incl. flatten(2).transpose(1,2).
//
// No im2col matrix is ever written to HBM.  The sample x[C,F,H,W] is described to the TMA unit as a 4-D tensor of
// 32-bit elements (one element = the (pw=0,1) bf16 pair of a patch row):
//     dim0 = w' (W' = W/2, contiguous)   dim1 = ph (2, stride W)   dim2 = rr = f*H'+h' (F*H', stride 2W)
//     dim3 = c  (C, stride F*H*W)
// One CTA owns an M-tile of R = floor(128 / W') whole patch rows (R*W' <= 128 tokens) and ALL of K = 4C <= 512:
//   warp 0    : TMA producer — per 16-channel k-block one box {W', 2, R, 16} into a raw staging ring, then the weight
//               tiles W[n*256 .., kb*64 ..] (128B swizzle) for this CTA's N tiles into a 3-stage ring whose third
//               stage is the raw staging area, free once the transform is done (weights stay in L2)
//   warps 2-5 : transform — thread = token row: gathers its 8 x 16 bytes {c, c+1} x {ph} x {pw} from the raw box and
//               writes them as the canonical K-major SWIZZLE_128B operand tile (what a 2-D TMA of an im2col matrix
//               would have produced); the A operand (<= 128 KB) then stays resident in smem for all N tiles
//   warp 1    : one thread issues tcgen05.mma M128 N256 K16, fp32 accumulators double-buffered in TMEM
//   warps 2-5 : afterwards the epilogue: tcgen05.ld -> + bias -> bf16 -> row (token - tok_lo) of `out`
// Tokens outside [tok_lo, tok_hi) are computed but not stored (sequence parallelism: a rank's token range need not
// start on a patch-row boundary).
#include "common.cuh"

namespace {

constexpr int BM = 128, BN = 256, BK = 64;
constexpr int A_BYTES = BM * BK * 2;   // 16 KB canonical A k-block
constexpr int W_BYTES = BN * BK * 2;   // 32 KB weight tile
constexpr int RAW_BYTES = 16 * 128 * 8;  // 16 channels x <=128 tokens x (2 ph x 2 pw x bf16) = 16 KB
constexpr int W_STAGES = 3, RAW_STAGES = 2, MAX_KB = 8;   // W stage 2 aliases the two raw staging buffers
constexpr int BAR_BYTES = 256;
constexpr int NUM_THREADS = 192;

struct PEParams {
  __nv_bfloat16* out;
  long long ldo;
  const __nv_bfloat16* bias;
  long long tok_lo, tok_hi;
  int row_lo;     // first patch row (f*H'+h') of tile 0
  int R, Wp;      // patch rows per tile, tokens per patch row
  int N, k_blocks, n_tiles;
  int n_per_cta;  // N tiles handled by one CTA (blockIdx.y selects the range)
};

__global__ void __launch_bounds__(NUM_THREADS, 1)
patch_embed_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW, const PEParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                                   // k_blocks x 16 KB
  uint8_t* sW = sA + p.k_blocks * A_BYTES;              // 2 x 32 KB (+ stage 2 = sRaw)
  uint8_t* sRaw = sW + (W_STAGES - 1) * W_BYTES;        // 2 x 16 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(sRaw + RAW_STAGES * RAW_BYTES);
  uint64_t* raw_full = bars;                 // [2]
  uint64_t* raw_empty = raw_full + RAW_STAGES;   // [2]
  uint64_t* a_ready = raw_empty + RAW_STAGES;    // [8]
  uint64_t* w_full = a_ready + MAX_KB;           // [3]
  uint64_t* w_empty = w_full + W_STAGES;         // [3]
  uint64_t* acc_full = w_empty + W_STAGES;       // [2]
  uint64_t* acc_empty = acc_full + 2;            // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < RAW_STAGES; ++i) {
      vc::mbar_init(&raw_full[i], 1);
      vc::mbar_init(&raw_empty[i], 128);
    }
    for (int i = 0; i < MAX_KB; ++i) vc::mbar_init(&a_ready[i], 128);
    for (int i = 0; i < W_STAGES; ++i) {
      vc::mbar_init(&w_full[i], 1);
      vc::mbar_init(&w_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      vc::mbar_init(&acc_full[i], 1);
      vc::mbar_init(&acc_empty[i], 128);
    }
    vc::fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    vc::tma_prefetch_desc(&tmX);
    vc::tma_prefetch_desc(&tmW);
  }
  if (warp == 2) vc::tmem_alloc(tmem_slot, 512);
  vc::tc_fence_before();
  __syncthreads();
  vc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int tile_row0 = p.row_lo + blockIdx.x * p.R;   // first patch row of this CTA
  const int tile_tokens = p.R * p.Wp;                  // valid A rows (<= 128)
  const uint32_t raw_tx = 16u * (uint32_t)tile_tokens * 8u;
  const int n_begin = blockIdx.y * p.n_per_cta;
  const int n_end = min(p.n_tiles, n_begin + p.n_per_cta);

  if (warp == 0) {
    // ------------------------------------------------ TMA producer
    if (vc::elect_one()) {
      for (int kb = 0; kb < p.k_blocks; ++kb) {
        const int s = kb % RAW_STAGES;
        vc::mbar_wait(&raw_empty[s], ((kb / RAW_STAGES) & 1) ^ 1);
        vc::mbar_expect_tx(&raw_full[s], raw_tx);
        vc::tma_load_4d(sRaw + s * RAW_BYTES, &tmX, &raw_full[s], 0, 0, tile_row0, kb * 16);
      }
      int stage = 0;
      uint32_t phase = 0;
      bool raw_free = false;
      for (int n = n_begin; n < n_end; ++n) {
        for (int kb = 0; kb < p.k_blocks; ++kb) {
          if (stage == W_STAGES - 1 && !raw_free) {   // stage 2 lives in the raw staging area: wait for the transform
            for (int i = 0; i < p.k_blocks; ++i) vc::mbar_wait(&a_ready[i], 0);
            raw_free = true;
          }
          vc::mbar_wait(&w_empty[stage], phase ^ 1);
          vc::mbar_expect_tx(&w_full[stage], W_BYTES);
          vc::tma_load_2d(sW + stage * W_BYTES, &tmW, &w_full[stage], kb * BK, n * BN);
          if (++stage == W_STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------ MMA issuer
    if (vc::elect_one()) {
      constexpr uint32_t idesc = vc::umma_idesc_bf16(BM, BN, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int n = n_begin; n < n_end; ++n) {
        vc::mbar_wait(&acc_empty[acc], acc_phase ^ 1);
        vc::tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < p.k_blocks; ++kb) {
          if (n == n_begin) vc::mbar_wait(&a_ready[kb], 0);   // A k-block transformed (stays resident afterwards)
          vc::mbar_wait(&w_full[stage], phase);
          vc::tc_fence_after();
          const uint32_t a0 = vc::smem_u32(sA + kb * A_BYTES);
          const uint32_t b0 = vc::smem_u32(sW + stage * W_BYTES);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            vc::umma_ss(d_tmem, vc::umma_desc_sw128(a0 + k * 32, 16, 1024), vc::umma_desc_sw128(b0 + k * 32, 16, 1024),
                        idesc, (kb | k) != 0 ? 1u : 0u);
          }
          vc::umma_commit(&w_empty[stage]);
          if (++stage == W_STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        vc::umma_commit(&acc_full[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else {
    // ------------------------------------------------ transform (implicit im2col), then epilogue
    const int quarter = warp & 3;              // TMEM lane quarter this warp may read
    const int m = quarter * 32 + lane;         // A row == accumulator row == token of the tile
    const bool m_ok = m < tile_tokens;
    const int r = m_ok ? m / p.Wp : 0;
    const int wq = m_ok ? m - r * p.Wp : 0;
    const int c_stride = p.R * 2 * p.Wp;       // raw box: [c][r][ph][w'] 32-bit elements
    const uint32_t sw = (uint32_t)(m & 7);
    for (int kb = 0; kb < p.k_blocks; ++kb) {
      const int s = kb % RAW_STAGES;
      vc::mbar_wait(&raw_full[s], (kb / RAW_STAGES) & 1);
      const uint32_t* raw = reinterpret_cast<const uint32_t*>(sRaw + s * RAW_BYTES) + (r * 2) * p.Wp + wq;
      uint8_t* arow = sA + kb * A_BYTES + (m >> 3) * 1024 + (m & 7) * 128;
#pragma unroll
      for (int j = 0; j < 8; ++j) {            // 16-byte chunk j = channels 2j, 2j+1 of this k-block
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (m_ok) {
          const uint32_t* c0 = raw + (2 * j) * c_stride;
          v.x = c0[0];                         // c   , ph 0, (pw 0, pw 1)
          v.y = c0[p.Wp];                      // c   , ph 1
          v.z = c0[c_stride];                  // c+1 , ph 0
          v.w = c0[c_stride + p.Wp];           // c+1 , ph 1
        }
        *reinterpret_cast<uint4*>(arow + (((uint32_t)j ^ sw) << 4)) = v;
      }
      vc::fence_proxy_async_smem();            // generic-proxy stores -> visible to tcgen05.mma's smem reads
      vc::mbar_arrive(&a_ready[kb]);
      vc::mbar_arrive(&raw_empty[s]);
    }

    const long long token = (long long)tile_row0 * p.Wp + m;
    const bool row_ok = m_ok && token >= p.tok_lo && token < p.tok_hi;
    __nv_bfloat16* crow = p.out + (row_ok ? token - p.tok_lo : 0) * p.ldo;
    int acc = 0;
    uint32_t acc_phase = 0;
    auto load_bias = [&](int n) {   // lane l holds bias[n*BN + 8l .. +8)
      const int col = n * BN + lane * 8;
      return col < p.N ? __ldg(reinterpret_cast<const uint4*>(p.bias + col)) : make_uint4(0u, 0u, 0u, 0u);
    };
    uint4 bias_cur = load_bias(n_begin);
    for (int n = n_begin; n < n_end; ++n) {
      const uint4 bias_nxt = n + 1 < n_end ? load_bias(n + 1) : make_uint4(0u, 0u, 0u, 0u);
      vc::mbar_wait(&acc_full[acc], acc_phase);
      vc::tc_fence_after();
      const uint32_t t_addr = tmem_base + acc * BN + (static_cast<uint32_t>(quarter * 32) << 16);
      // Round 2.  The ncu source page of the first form (tcgen05.ld, wait, bias load, use) put half of all stall samples on
      // the first use of the bias registers: the epilogue, not the tensor pipe (24 % active), set the pace.  Now
      //   * the tile's 256 bias values are fetched ONCE per tile, one 16-byte vector per lane, a whole tile ahead
      //     (bias_nxt below), and handed to the lanes that need them by warp shuffles: no global load in the chunk loop;
      //   * the TMEM load of chunk c+1 is issued before chunk c is converted and stored.
      uint32_t rg[2][32];
      vc::tmem_ld32(t_addr, rg[0]);
#pragma unroll
      for (int c = 0; c < BN / 32; ++c) {
        vc::tmem_ld_wait();
        if (c + 1 < BN / 32) vc::tmem_ld32(t_addr + (c + 1) * 32, rg[(c + 1) & 1]);
        const uint32_t(&r)[32] = rg[c & 1];
        const int col0 = n * BN + c * 32;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int src = c * 4 + g;   // lane holding columns [n*BN + src*8, +8)
          const uint32_t bx = __shfl_sync(0xffffffffu, bias_cur.x, src), by = __shfl_sync(0xffffffffu, bias_cur.y, src);
          const uint32_t bz = __shfl_sync(0xffffffffu, bias_cur.z, src), bw = __shfl_sync(0xffffffffu, bias_cur.w, src);
          const int col = col0 + g * 8;
          if (row_ok && col < p.N) {
            const float2 b0 = vc::unpack_bf16x2(bx), b1 = vc::unpack_bf16x2(by), b2 = vc::unpack_bf16x2(bz),
                         b3 = vc::unpack_bf16x2(bw);
            uint4 o;
            o.x = vc::pack_bf16x2(__uint_as_float(r[g * 8 + 0]) + b0.x, __uint_as_float(r[g * 8 + 1]) + b0.y);
            o.y = vc::pack_bf16x2(__uint_as_float(r[g * 8 + 2]) + b1.x, __uint_as_float(r[g * 8 + 3]) + b1.y);
            o.z = vc::pack_bf16x2(__uint_as_float(r[g * 8 + 4]) + b2.x, __uint_as_float(r[g * 8 + 5]) + b2.y);
            o.w = vc::pack_bf16x2(__uint_as_float(r[g * 8 + 6]) + b3.x, __uint_as_float(r[g * 8 + 7]) + b3.y);
            *reinterpret_cast<uint4*>(crow + col) = o;
          }
        }
      }
      bias_cur = bias_nxt;
      vc::tc_fence_before();
      vc::mbar_arrive(&acc_empty[acc]);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  }

  vc::tc_fence_before();
  __syncthreads();
  if (warp == 2) vc::tmem_dealloc(tmem_base, 512);
}

}  // namespace

extern "C" int vc_patch_embed_supported(int C, int F, int H, int W, int D) {
  if (C <= 0 || F <= 0 || H <= 0 || W <= 0 || D <= 0) return 0;
  if (C % 16 || 4 * C > MAX_KB * BK) return 0;          // whole 16-channel k-blocks, K <= 512 resident in smem
  if (H % 2 || W % 8 || W / 2 > BM) return 0;           // TMA: 16-byte strides (W*2 B) and a patch row per box
  if (D % 8) return 0;
  return 1;
}

extern "C" int vc_patch_embed(const void* x, int C, int F, int H, int W, const void* weight, long long ldw,
                              const void* bias, int D, void* out, long long ldo, long long tok_lo, long long tok_hi,
                              void* stream) {
  VC_CHECK_ARG(x && weight && bias && out, "vc_patch_embed: null operand");
  VC_CHECK_ARG(vc_patch_embed_supported(C, F, H, W, D),
               "vc_patch_embed: unsupported shape C=%d F=%d H=%d W=%d D=%d (need C %% 16 == 0, C <= 128, H even, "
               "W %% 8 == 0, W <= 256, D %% 8 == 0); use vc_patchify + vc_gemm_bf16",
               C, F, H, W, D);
  const int Hp = H / 2, Wp = W / 2;
  const long long L = (long long)F * Hp * Wp;
  VC_CHECK_ARG(tok_lo >= 0 && tok_lo <= tok_hi && tok_hi <= L, "vc_patch_embed: bad token range [%lld, %lld) of %lld",
               tok_lo, tok_hi, L);
  VC_CHECK_ARG(ldw % 8 == 0 && ldw >= 4 * C && ldo % 8 == 0 && ldo >= D,
               "vc_patch_embed: ldw/ldo must be multiples of 8 elements and cover the row");
  VC_CHECK_ARG((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(weight) | reinterpret_cast<uintptr_t>(out) |
                reinterpret_cast<uintptr_t>(bias)) % 16 == 0,
               "vc_patch_embed: operands must be 16-byte aligned");
  if (tok_hi == tok_lo) return 0;

  const int R = BM / Wp;
  const int row_lo = (int)(tok_lo / Wp);
  const int row_hi = (int)((tok_hi + Wp - 1) / Wp);
  const int m_tiles = (row_hi - row_lo + R - 1) / R;

  CUtensorMap tx, tw;
  {
    uint64_t dims[4] = {(uint64_t)Wp, 2, (uint64_t)F * Hp, (uint64_t)C};
    uint64_t strides[3] = {(uint64_t)W * 2, (uint64_t)W * 4, (uint64_t)F * H * W * 2};
    uint32_t box[4] = {(uint32_t)Wp, 2, (uint32_t)R, 16};
    int rc = vc_host::make_tmap(&tx, x, vc_host::TMAP_U32, vc_host::TMAP_SWIZZLE_NONE, 4, dims, strides, box);
    if (rc) return rc;
  }
  {
    uint64_t dims[2] = {(uint64_t)(4 * C), (uint64_t)D};
    uint64_t strides[1] = {(uint64_t)ldw * 2};
    uint32_t box[2] = {BK, BN};
    int rc = vc_host::make_tmap_bf16(&tw, weight, 2, dims, strides, box);
    if (rc) return rc;
  }
  PEParams p;
  p.out = (__nv_bfloat16*)out; p.ldo = ldo;
  p.bias = (const __nv_bfloat16*)bias;
  p.tok_lo = tok_lo; p.tok_hi = tok_hi;
  p.row_lo = row_lo; p.R = R; p.Wp = Wp;
  p.N = D; p.k_blocks = 4 * C / BK; p.n_tiles = (D + BN - 1) / BN;
  // N tiles are split over blockIdx.y when that fills the 148 SMs' waves better (cost model: one transform prologue
  // ~1.5 N-tile times per CTA, one CTA per SM)
  const int sms = vc_host::sm_count();
  int best_split = 1;
  double best_cost = 1e30;
  for (int split = 1; split <= 4 && split <= p.n_tiles; split *= 2) {
    const int per = (p.n_tiles + split - 1) / split;
    const long long ctas = (long long)m_tiles * ((p.n_tiles + per - 1) / per);
    const double cost = (double)((ctas + sms - 1) / sms) * (1.5 + per);
    if (cost < best_cost - 1e-9) {
      best_cost = cost;
      best_split = split;
    }
  }
  p.n_per_cta = (p.n_tiles + best_split - 1) / best_split;
  const int n_groups = (p.n_tiles + p.n_per_cta - 1) / p.n_per_cta;
  const int smem_bytes = p.k_blocks * A_BYTES + (W_STAGES - 1) * W_BYTES + RAW_STAGES * RAW_BYTES + BAR_BYTES + 1024;
  VC_ENSURE_DYN_SMEM(patch_embed_kernel, smem_bytes, "vc_patch_embed");
  patch_embed_kernel<<<dim3(m_tiles, n_groups), NUM_THREADS, smem_bytes, (cudaStream_t)stream>>>(tx, tw, p);
  VC_RETURN_LAUNCH_STATUS("vc_patch_embed");
}
