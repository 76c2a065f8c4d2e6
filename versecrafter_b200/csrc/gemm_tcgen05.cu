// bf16 GEMM  C[M,N] = epilogue(A[M,K] . W[N,K]^T)  on tcgen05 tensor cores (SURVEY.md §2.2 K8/K9/K10, K1/K2 after
// patchify).  Replaces every nn.Linear on the hot path (wan_transformer3d.py:366-369, :557-559; after_proj /
// before_proj wan_transformer3d_versecrafter.py:105-121; Conv3d patch embeddings :199-201, wan_transformer3d.py:758).
//
// Persistent, warp-specialised, one CTA per SM:
//   warp 0   : TMA producer  (A tile 128x64, W tile 256x64, 128B swizzle, 4-stage mbarrier ring)
//   warp 1   : one elected thread issues tcgen05.mma M=128 N=256 K=16 (4 per k-block), fp32 accumulators in TMEM
//   warp 2   : TMEM allocator (512 columns = 2 accumulator stages of 256) + epilogue
//   warps 2-9: epilogue (lane quarter = warp % 4, column half = (warp - 2) / 4): tcgen05.ld 32 columns at a time,
//              software-pipelined -> bias / GELU-tanh / gate / scale / residual -> bf16 global
// The accumulator is double-buffered in TMEM so tile i's epilogue overlaps tile i+1's main loop.
// Epilogue rounding points follow the reference's op-by-op bf16 arithmetic.
#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace {

constexpr int BM = 128, BN = 256, BK = 64, STAGES = 4;
constexpr int A_BYTES = BM * BK * 2;  // 16 KB
constexpr int B_BYTES = BN * BK * 2;  // 32 KB
constexpr int BAR_BYTES = 256;
constexpr int SMEM_BYTES = STAGES * (A_BYTES + B_BYTES) + BAR_BYTES + 1024;  // +1024 for manual alignment
constexpr int GROUP_M = 16;
constexpr int EPI_WARPS = 8;
constexpr int NUM_THREADS = 64 + 32 * EPI_WARPS;  // TMA warp, MMA warp, epilogue warps

enum : int { EPI_BIAS = 1, EPI_GELU = 2, EPI_GATE = 4, EPI_RES = 8, EPI_SCALE = 16, EPI_SCATTER = 32, SCHED_MULTICAST = 64,
             SCHED_PAIR = 128, SCHED_SINGLE = 256 };
constexpr int SCATTER_HEAD_DIM = 128;  // EPI_SCATTER: a head is 128 columns (256 bytes), so a 32-column chunk has one head

struct GemmParams {
  long long M, N, K;
  __nv_bfloat16* C;
  long long ldc;
  const __nv_bfloat16* bias;      // [N]
  const __nv_bfloat16* gate;      // [batch, N] rows gate_stride apart
  long long gate_stride;
  long long rows_per_batch;
  const __nv_bfloat16* residual;  // [M, N] rows ldr apart (may alias C)
  long long ldr;
  float scale;
  int m_tiles, n_tiles;
  int group_m;  // M tiles per rasterisation group (see vc_gemm_bf16)
  int l2_hint;  // bit 0: A loads evict_last; bit 1: W loads evict_first; bit 2: streaming C stores
  // EPI_SCATTER (vc_gemm_bf16_scatter): C is not one matrix but the head-scattered receive buffers of the peer GPUs
  __nv_bfloat16* peer[8];   // base of this matrix' block ([n_batch, l_total, heads_per_peer, 128]) in every peer's buffer
  int heads_per_peer;
  long long l_total, pos_offset;
};

__device__ __forceinline__ void tile_coords(int t, int m_tiles, int n_tiles, int group_m, int& m, int& n) {
  const int per_group = group_m * n_tiles;
  const int g = t / per_group;
  const int r = t - g * per_group;
  const int m0 = g * group_m;
  const int gm = min(group_m, m_tiles - m0);
  m = m0 + r % gm;
  n = r / gm;
}

__device__ __forceinline__ float gelu_tanh(float x) {
  // 0.5 x (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3)));  tanh(u) = 1 - 2 / (1 + exp(2u))
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  const float t = 1.0f - __fdividef(2.0f, 1.0f + __expf(2.0f * u));
  return 0.5f * x * (1.0f + t);
}

// Epilogue of one 128-row x 256-column accumulator tile, split over EPI_WARPS = 8 warps: warp w owns TMEM lane quarter
// w % 4 (= tile rows quarter*32 .. +31, one row per thread) and column half (w - 2) / 4 (128 columns = 4 chunks of 32).
// Per chunk: tcgen05.ld -> bias / GELU-tanh / gate / scale / residual (bf16 roundings as the reference's op-by-op
// arithmetic) -> bf16 16-byte stores; the next chunk's tcgen05.ld is in flight during the math.
// (Round 1 used 4 warps x 8 chunks with the load waited for immediately: at K = 1536 the epilogue of a tile then took
// longer than its 24 k-blocks of MMAs and the kernel fell to 0.71 of the bf16 peak, below cuBLASLt.)
template <int EPI>
__device__ __forceinline__ void epilogue_chunk(const GemmParams& p, const uint32_t* r, long long col0, bool row_ok,
                                               __nv_bfloat16* crow, const __nv_bfloat16* rrow, const __nv_bfloat16* grow,
                                               long long srow) {
  if (row_ok && col0 < p.N) {
    if (EPI & EPI_SCATTER) {
      // head-scatter store (Ulysses all-to-all inside the epilogue): columns [col0, col0+32) belong to ONE head; it lives
      // at the peer that owns the head, row (b, global token) of that peer's [n_batch, l_total, heads_per_peer, 128] block
      const int head = (int)(col0 / SCATTER_HEAD_DIM);
      const int peer = head / p.heads_per_peer;
      const int hl = head - peer * p.heads_per_peer;
      crow = p.peer[peer] + (srow + hl) * SCATTER_HEAD_DIM - (long long)head * SCATTER_HEAD_DIM;  // + col below
    }
#pragma unroll
    for (int g = 0; g < 4; ++g) {  // 8 columns = one 16-byte store
      const long long col = col0 + g * 8;
      if (col < p.N) {
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[g * 8 + j]);
        if (EPI & EPI_BIAS) {
          const uint4 bu = __ldg(reinterpret_cast<const uint4*>(p.bias + col));
          const float2 b0 = vc::unpack_bf16x2(bu.x), b1 = vc::unpack_bf16x2(bu.y), b2 = vc::unpack_bf16x2(bu.z),
                       b3 = vc::unpack_bf16x2(bu.w);
          v[0] += b0.x; v[1] += b0.y; v[2] += b1.x; v[3] += b1.y;
          v[4] += b2.x; v[5] += b2.y; v[6] += b3.x; v[7] += b3.y;
        }
        if (EPI & (EPI_GELU | EPI_GATE | EPI_RES | EPI_SCALE)) {
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = vc::bf16_round(v[j]);  // the Linear's own bf16 output
        }
        if (EPI & EPI_GELU) {
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = gelu_tanh(v[j]);
        }
        if (EPI & EPI_GATE) {
          const uint4 gu = __ldg(reinterpret_cast<const uint4*>(grow + col));
          const float2 g0 = vc::unpack_bf16x2(gu.x), g1 = vc::unpack_bf16x2(gu.y), g2 = vc::unpack_bf16x2(gu.z),
                       g3 = vc::unpack_bf16x2(gu.w);
          v[0] *= g0.x; v[1] *= g0.y; v[2] *= g1.x; v[3] *= g1.y;
          v[4] *= g2.x; v[5] *= g2.y; v[6] *= g3.x; v[7] *= g3.y;
        }
        if (EPI & EPI_SCALE) {
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] *= p.scale;
        }
        if (EPI & EPI_RES) {
          const uint4 ru = *reinterpret_cast<const uint4*>(rrow + col);
          const float2 r0 = vc::unpack_bf16x2(ru.x), r1 = vc::unpack_bf16x2(ru.y), r2 = vc::unpack_bf16x2(ru.z),
                       r3 = vc::unpack_bf16x2(ru.w);
          if (EPI & (EPI_GATE | EPI_SCALE)) {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = vc::bf16_round(v[j]);  // y*gate (or hint*scale) rounds first
          }
          v[0] += r0.x; v[1] += r0.y; v[2] += r1.x; v[3] += r1.y;
          v[4] += r2.x; v[5] += r2.y; v[6] += r3.x; v[7] += r3.y;
        }
        uint4 o;
        o.x = vc::pack_bf16x2(v[0], v[1]);
        o.y = vc::pack_bf16x2(v[2], v[3]);
        o.z = vc::pack_bf16x2(v[4], v[5]);
        o.w = vc::pack_bf16x2(v[6], v[7]);
        if (p.l2_hint & 4) __stcs(reinterpret_cast<uint4*>(crow + col), o);
        else *reinterpret_cast<uint4*>(crow + col) = o;
      }
    }
  }
}

template <int EPI>
__device__ __forceinline__ void epilogue_tile(const GemmParams& p, uint32_t t_addr, int m, int n, int quarter, int half,
                                              int lane) {
  const long long row = (long long)m * BM + quarter * 32 + lane;
  const bool row_ok = row < p.M;
  __nv_bfloat16* crow = p.C + row * p.ldc;
  const __nv_bfloat16* rrow = (EPI & EPI_RES) ? p.residual + row * p.ldr : nullptr;
  const __nv_bfloat16* grow =
      (EPI & EPI_GATE) ? p.gate + (row_ok ? row / p.rows_per_batch : 0) * p.gate_stride : nullptr;
  long long srow = 0;
  if (EPI & EPI_SCATTER) {  // (b * l_total + global token) * heads_per_peer: the row of local head 0 in a peer's block
    const long long b = row_ok ? row / p.rows_per_batch : 0;
    srow = (b * p.l_total + p.pos_offset + (row - b * p.rows_per_batch)) * p.heads_per_peer;
  }
  constexpr int CHUNKS = BN / 32 / 2;  // per warp
  const int c0 = half * CHUNKS;
  const long long colbase = (long long)n * BN + c0 * 32;
  uint32_t ra[32], rb[32];
  vc::tmem_ld32(t_addr + c0 * 32, ra);
#pragma unroll
  for (int c = 0; c < CHUNKS; c += 2) {
    vc::tmem_ld_wait();
    vc::tmem_ld32(t_addr + (c0 + c + 1) * 32, rb);
    epilogue_chunk<EPI>(p, ra, colbase + c * 32, row_ok, crow, rrow, grow, srow);
    vc::tmem_ld_wait();
    if (c + 2 < CHUNKS) vc::tmem_ld32(t_addr + (c0 + c + 2) * 32, ra);
    epilogue_chunk<EPI>(p, rb, colbase + (c + 1) * 32, row_ok, crow, rrow, grow, srow);
  }
}

// kMC: the grid is launched as clusters of 2 CTAs that work on two vertically adjacent M tiles of the same N tile; each
// CTA fetches its own A tile and HALF of the shared W tile, multicast into both CTAs' shared memory (L2 -> SM operand
// traffic 32 KB instead of 48 KB per k-block and CTA).  A stage is free only when BOTH consumers have released it (their
// tcgen05.commit arrives on both CTAs' `empty` barriers), because the peer's multicast writes into this CTA's stage.
template <int EPI, bool kMC>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * A_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * (A_BYTES + B_BYTES));
  uint64_t* empty = full + STAGES;
  uint64_t* acc_full = empty + STAGES;
  uint64_t* acc_empty = acc_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) {
      vc::mbar_init(&full[i], 1);
      vc::mbar_init(&empty[i], kMC ? 2 : 1);
    }
    for (int i = 0; i < 2; ++i) {
      vc::mbar_init(&acc_full[i], 1);
      vc::mbar_init(&acc_empty[i], 32 * EPI_WARPS);
    }
    vc::fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    vc::tma_prefetch_desc(&tmA);
    vc::tma_prefetch_desc(&tmB);
  }
  if (warp == 2) vc::tmem_alloc(tmem_slot, 512);
  vc::tc_fence_before();
  __syncthreads();
  vc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (kMC) vc::cluster_sync();  // the peer's barriers are initialised before anything is multicast to them

  // work items: tiles, or (kMC) pairs of M tiles handled by one cluster; both CTAs of a cluster walk the same list
  const int crank = kMC ? (int)vc::cluster_ctarank() : 0;
  const int t_first = kMC ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int t_step = kMC ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int sched_m = kMC ? (p.m_tiles + 1) / 2 : p.m_tiles;
  const int sched_g = kMC ? max(p.group_m / 2, 1) : p.group_m;
  const int num_tiles = sched_m * p.n_tiles;
  const int k_blocks = (int)((p.K + BK - 1) / BK);

  if (warp == 0) {
    // ------------------------------------------------ TMA producer
    if (vc::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      const uint64_t pol_a = (p.l2_hint & 1) ? vc::L2_EVICT_LAST : vc::L2_EVICT_NORMAL;
      const uint64_t pol_b = (p.l2_hint & 2) ? vc::L2_EVICT_FIRST : vc::L2_EVICT_NORMAL;
      for (int t = t_first; t < num_tiles; t += t_step) {
        int m, n;
        tile_coords(t, sched_m, p.n_tiles, sched_g, m, n);
        if (kMC) m = 2 * m + crank;  // may be one past the last M tile: loads are zero-filled, stores masked
        for (int kb = 0; kb < k_blocks; ++kb) {
          vc::mbar_wait(&empty[stage], phase ^ 1);
          vc::mbar_expect_tx(&full[stage], A_BYTES + B_BYTES);
          vc::tma_load_2d_hint(sA + stage * A_BYTES, &tmA, &full[stage], kb * BK, m * BM, pol_a);
          if (kMC)
            vc::tma_load_2d_mc(sB + stage * B_BYTES + crank * (B_BYTES / 2), &tmB, &full[stage], kb * BK,
                               n * BN + crank * (BN / 2), (uint16_t)3);
          else
            vc::tma_load_2d_hint(sB + stage * B_BYTES, &tmB, &full[stage], kb * BK, n * BN, pol_b);
          if (++stage == STAGES) {
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
      for (int t = t_first; t < num_tiles; t += t_step) {
        vc::mbar_wait(&acc_empty[acc], acc_phase ^ 1);
        vc::tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < k_blocks; ++kb) {
          vc::mbar_wait(&full[stage], phase);
          vc::tc_fence_after();
          const uint32_t a0 = vc::smem_u32(sA + stage * A_BYTES);
          const uint32_t b0 = vc::smem_u32(sB + stage * B_BYTES);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            vc::umma_ss(d_tmem, vc::umma_desc_sw128(a0 + k * 32, 16, 1024), vc::umma_desc_sw128(b0 + k * 32, 16, 1024),
                        idesc, (kb | k) != 0 ? 1u : 0u);
          }
          if (kMC) vc::umma_commit_mc(&empty[stage], (uint16_t)3);  // both CTAs' producers wait for both consumers
          else vc::umma_commit(&empty[stage]);                       // frees the smem stage when these MMAs retire
          if (++stage == STAGES) {
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
    // ------------------------------------------------ epilogue (warps 2..9 -> TMEM lane quarters 2,3,0,1,2,3,0,1)
    const int quarter = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = t_first; t < num_tiles; t += t_step) {
      int m, n;
      tile_coords(t, sched_m, p.n_tiles, sched_g, m, n);
        if (kMC) m = 2 * m + crank;  // may be one past the last M tile: loads are zero-filled, stores masked
      vc::mbar_wait(&acc_full[acc], acc_phase);
      vc::tc_fence_after();
      epilogue_tile<EPI>(p, tmem_base + acc * BN + (static_cast<uint32_t>(quarter * 32) << 16), m, n, quarter, (warp - 2) >> 2,
                         lane);
      vc::tc_fence_before();
      vc::mbar_arrive(&acc_empty[acc]);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  }

  vc::tc_fence_before();
  __syncthreads();
  if (kMC) vc::cluster_sync();  // the peer may still multicast into this CTA's smem / arrive on its barriers
  if (warp == 2) vc::tmem_dealloc(tmem_base, 512);
}

// ------------------------------------------------------------------------------------------------------------------------
// CTA-pair variant (`tcgen05.mma.cta_group::2`, M = 256 per pair).  A cluster of two CTAs (the two SMs of one TPC) owns a
// 256 x 256 output tile: CTA r holds rows [r*128, r*128+128) of the A tile and rows [r*128, ...) of the W tile (HALF of
// it) per k-block, 32 KB per stage instead of 48 KB, so the ring is 6 deep; each SM's tensor core reads its own A rows and
// both halves of W (the peer's half over the pair's shared-memory path) — the operand bytes each SM pulls out of shared
// memory and out of L2 per flop drop by a third, which is the energy-per-flop lever on a power-capped chip.
// Roles per CTA are those of the kernel above; only the leader (cluster rank 0) runs the MMA issuer.  Barriers:
//   full[s]      leader's only: expect_tx of BOTH CTAs' bytes (64 KB); every TMA load of either CTA credits it
//   empty[s]     one per CTA, count 1: the leader's tcgen05.commit multicasts the arrive to both producers
//   acc_full[a]  one per CTA, count 1: multicast commit at the end of a tile's k loop -> both epilogues
//   acc_empty[a] leader's only, count 512: the 256 epilogue threads of each CTA arrive (the peer's remotely)
constexpr int STAGES2 = 6;
constexpr int B2_BYTES = (BN / 2) * BK * 2;  // 16 KB: this CTA's half of the W tile
constexpr int SMEM2_BYTES = STAGES2 * (A_BYTES + B2_BYTES) + BAR_BYTES + 1024;

template <int EPI>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_bf16_2cta_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES2 * A_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES2 * (A_BYTES + B2_BYTES));
  uint64_t* empty = full + STAGES2;
  uint64_t* acc_full = empty + STAGES2;
  uint64_t* acc_empty = acc_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int crank = (int)vc::cluster_ctarank();

  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES2; ++i) {
      vc::mbar_init(&full[i], 1);
      vc::mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      vc::mbar_init(&acc_full[i], 1);
      vc::mbar_init(&acc_empty[i], 2 * 32 * EPI_WARPS);
    }
    vc::fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    vc::tma_prefetch_desc(&tmA);
    vc::tma_prefetch_desc(&tmB);
  }
  if (warp == 2) vc::tmem_alloc_2cta(tmem_slot, 512);
  vc::tc_fence_before();
  __syncthreads();
  vc::cluster_sync();  // both CTAs' barriers are initialised and both TMEM halves allocated before any cross-CTA signal
  vc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int t_first = (int)(blockIdx.x >> 1), t_step = (int)(gridDim.x >> 1);
  const int sched_m = (p.m_tiles + 1) / 2;
  const int sched_g = max(p.group_m / 2, 1);
  const int num_tiles = sched_m * p.n_tiles;
  const int k_blocks = (int)((p.K + BK - 1) / BK);

  if (warp == 0) {
    // ------------------------------------------------ TMA producer (both CTAs: own A rows, own half of W)
    if (vc::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      const uint64_t pol_a = (p.l2_hint & 1) ? vc::L2_EVICT_LAST : vc::L2_EVICT_NORMAL;
      const uint64_t pol_b = (p.l2_hint & 2) ? vc::L2_EVICT_FIRST : vc::L2_EVICT_NORMAL;
      for (int t = t_first; t < num_tiles; t += t_step) {
        int mp, n;
        tile_coords(t, sched_m, p.n_tiles, sched_g, mp, n);
        const int m = 2 * mp + crank;  // may be one past the last M tile: loads are zero-filled, stores masked
        for (int kb = 0; kb < k_blocks; ++kb) {
          vc::mbar_wait(&empty[stage], phase ^ 1);
          if (crank == 0) vc::mbar_expect_tx(&full[stage], 2 * (A_BYTES + B2_BYTES));
          vc::tma_load_2d_2cta(sA + stage * A_BYTES, &tmA, &full[stage], kb * BK, m * BM, pol_a);
          vc::tma_load_2d_2cta(sB + stage * B2_BYTES, &tmB, &full[stage], kb * BK, n * BN + crank * (BN / 2), pol_b);
          if (++stage == STAGES2) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------ MMA issuer (leader CTA only): M = 256, N = 256, K = 16
    if (crank == 0 && vc::elect_one()) {
      constexpr uint32_t idesc = vc::umma_idesc_bf16(2 * BM, BN, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = t_first; t < num_tiles; t += t_step) {
        vc::mbar_wait_cluster(&acc_empty[acc], acc_phase ^ 1);
        vc::tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < k_blocks; ++kb) {
          vc::mbar_wait_cluster(&full[stage], phase);
          vc::tc_fence_after();
          const uint32_t a0 = vc::smem_u32(sA + stage * A_BYTES);
          const uint32_t b0 = vc::smem_u32(sB + stage * B2_BYTES);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            vc::umma_ss_2cta(d_tmem, vc::umma_desc_sw128(a0 + k * 32, 16, 1024), vc::umma_desc_sw128(b0 + k * 32, 16, 1024),
                             idesc, (kb | k) != 0 ? 1u : 0u);
          }
          vc::umma_commit_2cta_mc(&empty[stage], (uint16_t)3);  // frees this stage in both CTAs when the MMAs retire
          if (++stage == STAGES2) {
            stage = 0;
            phase ^= 1;
          }
        }
        vc::umma_commit_2cta_mc(&acc_full[acc], (uint16_t)3);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else {
    // ------------------------------------------------ epilogue (warps 2..9 of both CTAs, each CTA its own 128 rows)
    const int quarter = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = t_first; t < num_tiles; t += t_step) {
      int mp, n;
      tile_coords(t, sched_m, p.n_tiles, sched_g, mp, n);
      const int m = 2 * mp + crank;
      vc::mbar_wait_cluster(&acc_full[acc], acc_phase);
      vc::tc_fence_after();
      epilogue_tile<EPI>(p, tmem_base + acc * BN + (static_cast<uint32_t>(quarter * 32) << 16), m, n, quarter, (warp - 2) >> 2,
                         lane);
      vc::tc_fence_before();
      vc::mbar_arrive_leader(&acc_empty[acc]);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  }

  vc::tc_fence_before();
  __syncthreads();
  vc::cluster_sync();  // the leader's MMAs read the peer's shared memory and signal its barriers until the very end
  if (warp == 2) vc::tmem_dealloc_2cta(tmem_base, 512);
}

template <int EPI>
int launch_2cta(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid, cudaStream_t s) {
  VC_ENSURE_DYN_SMEM((gemm_bf16_2cta_kernel<EPI>), SMEM2_BYTES, "vc_gemm_bf16");
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(NUM_THREADS);
  cfg.dynamicSmemBytes = SMEM2_BYTES;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_bf16_2cta_kernel<EPI>, ta, tb, p);
  if (e != cudaSuccess) {
    vc_host::set_error("vc_gemm_bf16: CTA-pair launch failed: %s", cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

template <int EPI, bool kMC>
int launch_impl(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid, cudaStream_t s) {
  VC_ENSURE_DYN_SMEM((gemm_bf16_kernel<EPI, kMC>), SMEM_BYTES, "vc_gemm_bf16");
  if (!kMC) {
    gemm_bf16_kernel<EPI, kMC><<<grid, NUM_THREADS, SMEM_BYTES, s>>>(ta, tb, p);
    return 0;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(NUM_THREADS);
  cfg.dynamicSmemBytes = SMEM_BYTES;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_bf16_kernel<EPI, kMC>, ta, tb, p);
  if (e != cudaSuccess) {
    vc_host::set_error("vc_gemm_bf16: cluster launch failed: %s", cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

// sched: 0 = one CTA per tile, 1 = cluster of 2 sharing W by TMA multicast, 2 = CTA pair (cta_group::2, M = 256)
template <int EPI>
int launch(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid, int sched, cudaStream_t s) {
  if (sched == 2) return launch_2cta<EPI>(ta, tb, p, grid, s);
  return sched == 1 ? launch_impl<EPI, true>(ta, tb, p, grid, s) : launch_impl<EPI, false>(ta, tb, p, grid, s);
}

using vc_host::sm_count;

}  // namespace

// `sc` != nullptr: the head-scatter epilogue (vc_gemm_bf16_scatter); its peer / layout fields are copied into the kernel
// parameters and C is unused.
static int gemm_entry(const void* A, long long lda, const void* W, long long ldw, void* C, long long ldc,
                      long long M, long long N, long long K, int epilogue, const void* bias, const void* gate,
                      long long gate_batch_stride, long long rows_per_batch, const void* residual,
                      long long ldr, float scale, void* stream, const GemmParams* sc) {
  if (sc) {  // placeholders that satisfy the dense-C checks below; no thread dereferences C with EPI_SCATTER
    C = const_cast<void*>(static_cast<const void*>(sc->peer[0]));
    ldc = N;
  }
  VC_CHECK_ARG(A && W && C, "vc_gemm_bf16: null operand");
  VC_CHECK_ARG(M >= 0 && N > 0 && K > 0, "vc_gemm_bf16: bad shape M=%lld N=%lld K=%lld", M, N, K);
  VC_CHECK_ARG(N % 8 == 0 && K % 8 == 0, "vc_gemm_bf16: N and K must be multiples of 8 (N=%lld K=%lld)", N, K);
  VC_CHECK_ARG(lda % 8 == 0 && ldw % 8 == 0 && ldc % 8 == 0 && lda >= K && ldw >= K && ldc >= N,
               "vc_gemm_bf16: leading dimensions must be multiples of 8 elements and cover the row");
  VC_CHECK_ARG((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(W) | reinterpret_cast<uintptr_t>(C)) % 16 == 0,
               "vc_gemm_bf16: operands must be 16-byte aligned");
  VC_CHECK_ARG(!(epilogue & EPI_BIAS) || bias, "vc_gemm_bf16: bias epilogue without bias");
  VC_CHECK_ARG(!(epilogue & EPI_GATE) || (gate && rows_per_batch > 0), "vc_gemm_bf16: gate epilogue without gate");
  VC_CHECK_ARG(!(epilogue & EPI_RES) || (residual && ldr % 8 == 0 && ldr >= N), "vc_gemm_bf16: residual epilogue without residual");
  if (M == 0) return 0;
  if (!sc) {
    // The kernel is persistent and multi-tile: tile (m, n) stores its C block while other CTAs / later waves still load
    // A[m rows, 0:K] for the other N tiles, so A must not overlap C (only `residual` may alias C: each element is read
    // and written by the same thread).  Interleaved column blocks of one row-strided buffer (same ld, disjoint
    // columns) are fine.
    const uintptr_t a0 = reinterpret_cast<uintptr_t>(A), c0 = reinterpret_cast<uintptr_t>(C);
    const uintptr_t a1 = a0 + (uintptr_t)((M - 1) * lda + K) * 2, c1 = c0 + (uintptr_t)((M - 1) * ldc + N) * 2;
    bool clash = a0 < c1 && c0 < a1;
    if (clash && lda == ldc) {
      const uintptr_t row_bytes = (uintptr_t)lda * 2;
      const uintptr_t lo = a0 < c0 ? a0 : c0;
      const uintptr_t ca = (a0 - lo) % row_bytes, cc = (c0 - lo) % row_bytes;  // column offsets within a row
      const bool a_fits = ca + (uintptr_t)K * 2 <= row_bytes, c_fits = cc + (uintptr_t)N * 2 <= row_bytes;
      if (a_fits && c_fits && (ca + (uintptr_t)K * 2 <= cc || cc + (uintptr_t)N * 2 <= ca)) clash = false;
    }
    VC_CHECK_ARG(!clash, "vc_gemm_bf16: A and C overlap (the persistent kernel re-reads A rows after C tiles of the same "
                         "rows are stored); write C to a separate buffer");
  }

  // VC_GEMM_MC=1 or the VC_GEMM_SCHED_MULTICAST bit: clusters of 2 CTAs sharing the W tile by TMA multicast (see the
  // kernel); needs >= 2 M tiles.  Same arithmetic, bit-identical results.
  // Schedule (same arithmetic, bit-identical results): 0 = one CTA per 128x256 tile; 1 = cluster of 2 sharing the W tile by
  // TMA multicast; 2 = CTA pair, one cta_group::2 MMA of M = 256 per pair.  Chosen by the VC_GEMM_SCHED_* bits of
  // `epilogue`, else VC_GEMM_SCHED = single | mc | pair (VC_GEMM_MC=1 is the older spelling of mc), else the default.
  static const int sched_env = []() {
    const char* e = getenv("VC_GEMM_SCHED");
    if (e && !strcmp(e, "single")) return 0;
    if (e && !strcmp(e, "mc")) return 1;
    if (e && !strcmp(e, "pair")) return 2;
    const char* m = getenv("VC_GEMM_MC");
    return (m && m[0] == '1') ? 1 : -1;
  }();
  // Default per shape, from same-box measurements on B200 (profiles/gemm_sched_r02c.jsonl alone; profiles/bench_c3_r02b_abab_*,
  // bench_c2_r02c_*, gpurun N=2 ABAB in the step): the CTA pair is faster alone at every c3 / c2 shape but one (+4..+15 %,
  // +11..+15 % at the M = 8190 rows of an 8-GPU rank) and wins inside the step at c2 (743 vs 758 ms) and under sequence
  // parallelism (2 GPUs: 2135 vs 2152 ms); inside the single-GPU c3 step (M = 65520, K >= 5120) the chip sits on its power
  // cap, the pair's GEMMs gain 5 % and self-attention loses 5 % of its clock, and the step is 1 % slower (4280 vs 4237 ms).
  const int sched_auto = (M > 40000 && K >= 4096) ? 0 : 2;
  int sched = (epilogue & SCHED_SINGLE) ? 0 : (epilogue & SCHED_PAIR) ? 2 : (epilogue & SCHED_MULTICAST) ? 1
              : sched_env >= 0 ? sched_env : sched_auto;
  if (M <= BM) sched = 0;  // the paired schedules need two M tiles
  const bool mc = sched != 0;
  CUtensorMap ta, tb;
  {
    uint64_t dims[2] = {(uint64_t)K, (uint64_t)M};
    uint64_t strides[1] = {(uint64_t)lda * 2};
    uint32_t box[2] = {BK, BM};
    int rc = vc_host::make_tmap_bf16(&ta, A, 2, dims, strides, box);
    if (rc) return rc;
  }
  {
    uint64_t dims[2] = {(uint64_t)K, (uint64_t)N};
    uint64_t strides[1] = {(uint64_t)ldw * 2};
    uint32_t box[2] = {BK, (uint32_t)(mc ? BN / 2 : BN)};  // multicast: each CTA of a pair fetches half of the W tile
    int rc = vc_host::make_tmap_bf16(&tb, W, 2, dims, strides, box);
    if (rc) return rc;
  }
  GemmParams p;
  p.M = M; p.N = N; p.K = K;
  p.C = (__nv_bfloat16*)C; p.ldc = ldc;
  p.bias = (const __nv_bfloat16*)bias;
  p.gate = (const __nv_bfloat16*)gate; p.gate_stride = gate_batch_stride;
  p.rows_per_batch = rows_per_batch > 0 ? rows_per_batch : 1;
  p.residual = (const __nv_bfloat16*)residual; p.ldr = ldr;
  p.scale = scale;
  for (int i = 0; i < 8; ++i) p.peer[i] = sc ? sc->peer[i] : nullptr;
  p.heads_per_peer = sc ? sc->heads_per_peer : 1;
  p.l_total = sc ? sc->l_total : 0;
  p.pos_offset = sc ? sc->pos_offset : 0;
  p.m_tiles = (int)((M + BM - 1) / BM);
  p.n_tiles = (int)((N + BN - 1) / BN);
  // Rasterisation: tiles are visited group by group, a group = group_m M-tiles x all N-tiles (M fastest), so the A rows of
  // a group are re-used from L2 by consecutive waves while the weight streams past once per group: DRAM reads are about
  // |A| + |W| * m_tiles / group_m as long as the A panel survives in L2 between waves.  Measured on B200 (ncu
  // dram__bytes_read per launch, M=65520, profiles/gemm_raster_r01.json): N=15360,K=5120 reads 5.6 GB at group_m=16,
  // 4.3 GB at 24, 6.2 GB at 33 and 10.5 GB at 48 (panel evicted; the c3 step gets 2.7 % slower: DRAM traffic costs power
  // under the 1 kW cap).  Loading A with an L2 evict_last policy keeps the taller panel resident: 3.8 GB at group_m=32
  // (W evict_first is harmful: it destroys the sharing of a W tile between the CTAs of one wave, 9.5 GB).  For
  // K=13824 no height in 8..18 gets below 10.4 GB (one wave's footprint exceeds what L2 retains).
  // A same-box A/B of the whole c3 step (two runs each, alternating) decided the default: group_m=16 without hints
  // 4191.9 / 4198.3 ms (GEMMs 1,340 / 1,334 TFLOP/s) vs evict_last + group_m=32 4200.2 / 4208.2 ms (1,325 / 1,322):
  // the lower DRAM traffic does not pay for itself, so the plain setting stays.  VC_GEMM_GROUP_M / VC_GEMM_L2_HINT
  // (bit 0: A evict_last, bit 1: W evict_first, bit 2: streaming C stores) remain for experiments.
  static const int group_m_env = []() {
    const char* e = getenv("VC_GEMM_GROUP_M");
    const int g = e ? atoi(e) : 0;
    return g >= 1 && g <= 512 ? g : 0;
  }();
  static const int l2_hint = []() { const char* e = getenv("VC_GEMM_L2_HINT"); return e ? atoi(e) & 7 : 0; }();
  p.group_m = group_m_env ? group_m_env : GROUP_M;
  p.l2_hint = l2_hint;
  const long long tiles = (long long)p.m_tiles * p.n_tiles;
  int grid = (int)(tiles < sm_count() ? tiles : sm_count());
  if (mc) {
    const long long pairs = (long long)((p.m_tiles + 1) / 2) * p.n_tiles;
    const int max_clusters = sm_count() / 2;
    grid = 2 * (int)(pairs < max_clusters ? pairs : max_clusters);
  }
  cudaStream_t s = (cudaStream_t)stream;
  int rc;
  switch (epilogue & ~(SCHED_MULTICAST | SCHED_PAIR | SCHED_SINGLE)) {
    case 0: rc = launch<0>(ta, tb, p, grid, sched, s); break;
    case EPI_BIAS: rc = launch<EPI_BIAS>(ta, tb, p, grid, sched, s); break;
    case EPI_BIAS | EPI_GELU: rc = launch<EPI_BIAS | EPI_GELU>(ta, tb, p, grid, sched, s); break;
    case EPI_BIAS | EPI_RES: rc = launch<EPI_BIAS | EPI_RES>(ta, tb, p, grid, sched, s); break;
    case EPI_BIAS | EPI_GATE | EPI_RES: rc = launch<EPI_BIAS | EPI_GATE | EPI_RES>(ta, tb, p, grid, sched, s); break;
    case EPI_BIAS | EPI_SCALE | EPI_RES: rc = launch<EPI_BIAS | EPI_SCALE | EPI_RES>(ta, tb, p, grid, sched, s); break;
    case EPI_BIAS | EPI_SCATTER:
      if (sc) { rc = launch<EPI_BIAS | EPI_SCATTER>(ta, tb, p, grid, sched, s); break; }
      [[fallthrough]];
    default:
      vc_host::set_error("vc_gemm_bf16: unsupported epilogue mask %d", epilogue);
      return -1;
  }
  if (rc) return rc;
  VC_RETURN_LAUNCH_STATUS("vc_gemm_bf16");
}

extern "C" int vc_gemm_bf16(const void* A, long long lda, const void* W, long long ldw, void* C, long long ldc,
                            long long M, long long N, long long K, int epilogue, const void* bias, const void* gate,
                            long long gate_batch_stride, long long rows_per_batch, const void* residual,
                            long long ldr, float scale, void* stream) {
  VC_CHECK_ARG(!(epilogue & EPI_SCATTER), "vc_gemm_bf16: the scatter epilogue is vc_gemm_bf16_scatter");
  return gemm_entry(A, lda, W, ldw, C, ldc, M, N, K, epilogue, bias, gate, gate_batch_stride, rows_per_batch, residual, ldr,
                    scale, stream, nullptr);
}

// Linear + Ulysses all-to-all in one kernel: Y = A . W^T + bias for this rank's token rows, every head's 128 columns
// stored straight into the receive buffer of the peer GPU that owns the head (same destination formula as
// vc_rmsnorm_rope_qkv_scatter; `peer_recv[i]` = base of this matrix' [n_batch, l_total, heads_per_peer, 128] block in
// peer i's buffer).  Used for v, which needs no normalisation: its exchange leaves the critical path and rides under
// the GEMM's own main loop (wan_transformer3d.py:383-389 `v = self.v(x)` + xfuser's SeqAllToAll4D).
extern "C" int vc_gemm_bf16_scatter(const void* A, long long lda, const void* W, long long ldw, long long M, long long N,
                                    long long K, const void* bias, long long rows_per_batch, long long pos_offset,
                                    void* const* peer_recv, int n_peers, int heads_per_peer, int head_dim,
                                    long long l_total, int sched, void* stream) {
  VC_CHECK_ARG(peer_recv && bias, "vc_gemm_bf16_scatter: null pointer");
  VC_CHECK_ARG(n_peers >= 1 && n_peers <= 8, "vc_gemm_bf16_scatter: n_peers=%d must be in 1..8", n_peers);
  VC_CHECK_ARG(head_dim == SCATTER_HEAD_DIM, "vc_gemm_bf16_scatter: head_dim=%d must be 128", head_dim);
  VC_CHECK_ARG(heads_per_peer > 0 && (long long)heads_per_peer * n_peers * head_dim == N,
               "vc_gemm_bf16_scatter: heads_per_peer * n_peers * head_dim must equal N=%lld", N);
  VC_CHECK_ARG(M >= 0 && rows_per_batch > 0 && M % rows_per_batch == 0 && pos_offset >= 0 &&
                   pos_offset + rows_per_batch <= l_total,
               "vc_gemm_bf16_scatter: bad rows / positions");
  VC_CHECK_ARG(sched == 0 || sched == SCHED_MULTICAST || sched == SCHED_PAIR || sched == SCHED_SINGLE,
               "vc_gemm_bf16_scatter: sched must be 0 or one VC_GEMM_SCHED_* bit");
  GemmParams sc = {};
  for (int i = 0; i < n_peers; ++i) {
    VC_CHECK_ARG(peer_recv[i] && (uintptr_t)peer_recv[i] % 16 == 0, "vc_gemm_bf16_scatter: bad peer pointer %d", i);
    sc.peer[i] = (__nv_bfloat16*)peer_recv[i];
  }
  sc.heads_per_peer = heads_per_peer;
  sc.l_total = l_total;
  sc.pos_offset = pos_offset;
  return gemm_entry(A, lda, W, ldw, nullptr, 0, M, N, K, EPI_BIAS | EPI_SCATTER | sched, bias, nullptr, 0, rows_per_batch,
                    nullptr, 0, 1.0f, stream, &sc);
}
