/* vc_b200.h — C ABI of libvc_b200.so: hand-written sm_100a (B200) kernels for VerseCrafter's denoise step.
 *
 * Conventions: raw device pointers, explicit sizes and strides (in ELEMENTS unless noted), the
 * caller's cudaStream_t passed as void*; every function only enqueues work on that stream — no allocation, no
 * synchronisation, no C++ exceptions.  Return value: 0 = OK, < 0 = invalid argument / unsupported (message in
 * vc_last_error()), > 0 = CUDA error code (cudaError_t, or 1000 + CUresult from tensor-map encoding).
 * All tensors are bf16 (uint16 storage) unless noted.  The reference has no native code; each entry point replaces
 * the PyTorch/library calls at the cited lines of /root/reference/versecrafter/models/.
 */
#ifndef VC_B200_H
#define VC_B200_H
#ifdef __cplusplus
extern "C" {
#endif

/* library / device probes */
int vc_version(void);            /* 100 = 0.1.0 */
int vc_build_arch(void);         /* 100: compiled for sm_100a only */
int vc_device_ok(void);          /* 0 if the current CUDA device is compute capability 10.x */
const char* vc_last_error(void); /* thread-local message of the last non-zero return */

/* K3/K4 — LayerNorm (fp32 statistics, eps inside rsqrt, no learned affine) fused with the AdaLN modulation.
 * mode 0: out[r,:] = LN(x[r,:]) * (1 + scale[b,:]) + shift[b,:],  b = r / rows_per_batch,
 *         shift/scale rows are mod_batch_stride elements apart  (WanLayerNorm + modulate,
 *         wan_transformer3d.py:326-346 with :591-592, :603-604; Head :641-643)
 * mode 1: out[r,:] = LN(x[r,:]) * gamma + beta   (norm3 affine, wan_transformer3d.py:548-550 used at :600);
 *         pass beta as shift_or_beta and gamma as scale_or_gamma.
 * D % 8 == 0, D <= 8192; ldx/ldo = row strides.  Schedule bits (same arithmetic, bit-identical results): mode | 16 forces
 * the register-resident kernel, mode | 32 the asynchronous row ring (cp.async.bulk prefetch, D > 1024); default: the ring for
 * rows of >= 4 KB (D >= 2048), registers below. */
int vc_ln_modulate(const void* x, long long ldx, void* out, long long ldo, const void* shift_or_beta,
                   const void* scale_or_gamma, long long mod_batch_stride, long long rows, long long rows_per_batch,
                   int D, float eps, int mode, void* stream);

/* K5 — in place on rows [q(D) | k(D) | ...] (row stride ld): the first n_mats (1 or 2) D-wide column blocks
 * <- RoPE(RMSNorm_D(.) * gain)  (block 0 uses gain_q, block 1 gain_k; the rest of the row, e.g. v, is untouched).
 * RMSNorm is over the FULL hidden dim D (WanRMSNorm, wan_transformer3d.py:307-323, applied before the head split
 * :385-386); RoPE is the 3-axis adjacent-pair rotation of rope_apply (:143-172) with per-axis (cos,sin) tables
 * rope_table = float2 [grid_f*nf | grid_h*nh | grid_w*nw] (fp32, built from fp64 angles), nf+nh+nw = head_dim/2.
 * Token position = pos_offset + (row % rows_per_batch)  (pos_offset = first token of this rank under sequence
 * parallelism); positions >= grid_f*grid_h*grid_w are normalised but not rotated.  apply_rope = 0: RMSNorm only
 * (cross-attention: q rows with n_mats = 1, :420; text-context [k | v] rows with n_mats = 1, :421).
 * apply_rope | 16 / | 32: schedule bits as in vc_ln_modulate (default here: the register-resident kernel). */
int vc_rmsnorm_rope_qk(void* qkv, long long ld, long long rows, long long rows_per_batch, long long pos_offset, int D,
                       int n_mats, const void* gain_q, const void* gain_k, const void* rope_table, int grid_f,
                       int grid_h, int grid_w, int nf, int nh, int nw, float eps, int apply_rope, void* stream);

/* K8/K9/K10 (+K1/K2 after vc_patchify) — C[M,N] = epilogue(A[M,K] . W[N,K]^T), fp32 accumulation on tcgen05.
 * W is an nn.Linear weight as stored ([out, in], row-major).  epilogue is a bit mask applied in this order, with a
 * bf16 rounding wherever the reference's op-by-op bf16 arithmetic has one:
 *   VC_EPI_BIAS  (1): + bias[N]                         (every nn.Linear on the path, wan_transformer3d.py:366-369)
 *   VC_EPI_GELU  (2): GELU-tanh                         (ffn.1, :558)
 *   VC_EPI_GATE  (4): * gate[b, N], b = row / rows_per_batch, gate rows gate_batch_stride apart  (:595, :607)
 *   VC_EPI_SCALE (16): * scale                          (hint * context_scale, wan_transformer3d_versecrafter.py:147)
 *   VC_EPI_RES   (8): + residual[M,N] (row stride ldr; may alias C)   (:595, :600, :607; before_proj + x :114)
 * Supported masks: 0, 1, 3, 9, 13, 25.  N % 8 == 0, K % 8 == 0, lda/ldw/ldc/ldr % 8 == 0, 16-byte aligned bases. */
#define VC_EPI_BIAS 1
#define VC_EPI_GELU 2
#define VC_EPI_GATE 4
#define VC_EPI_RES 8
#define VC_EPI_SCALE 16
/* schedule bit, not arithmetic: run as clusters of two CTAs that share the W tile by TMA multicast (same results) */
#define VC_GEMM_SCHED_MULTICAST 64
/* CTA pair: one tcgen05.mma.cta_group::2 of M = 256 per pair of CTAs (each holds its 128 A rows and half of the W tile) */
#define VC_GEMM_SCHED_PAIR 128
/* one CTA per 128 x 256 tile, whatever VC_GEMM_SCHED says */
#define VC_GEMM_SCHED_SINGLE 256
int vc_gemm_bf16(const void* A, long long lda, const void* W, long long ldw, void* C, long long ldc, long long M,
                 long long N, long long K, int epilogue, const void* bias, const void* gate,
                 long long gate_batch_stride, long long rows_per_batch, const void* residual, long long ldr,
                 float scale, void* stream);

/* a4 — the fp32 timestep-embedding island, wan_transformer3d_versecrafter.py:347-354:
 *   s  = [cos(t w_j) | sin(t w_j)], w_j = 10000^(-j/(freq_dim/2))  in fp64, rounded to fp32   (wan_transformer3d.py:39-49)
 *   e  = W2 silu(W0 s + b0) + b2                                    (time_embedding, :764-765)
 *   e0 = Wp silu(e) + bp                                            (time_projection, :766), viewed [B, 6, D]
 * fp32 arithmetic on the bf16-valued weights (w0 [D, freq_dim], w2 [D, D], wp [6D, D], row-major, contiguous), outputs
 * rounded once to bf16: e_out [B, D], e0_out [B, 6D].  t: fp64 [B] on the device (1 <= B <= 8).
 * workspace: caller-provided fp32 scratch of at least B * (freq_dim + 2 D) elements, 16-byte aligned. */
int vc_time_embed(const double* t, int B, int freq_dim, int D, const void* w0, const void* b0, const void* w2,
                  const void* b2, const void* wp, const void* bp, void* e_out, void* e0_out, float* workspace,
                  void* stream);

/* vc_time_embed without the workspace argument.  No counterpart in the reference: it exists for hosts that cannot allocate
 * device memory when they need the embedding — a loop whose memory pool was carved up before the run and is now closed,
 * code running on a thread or in a callback that must not touch the allocator, or a C caller that was itself only handed
 * input and output pointers.  The three intermediate vectors live in a fixed array compiled into the library
 * (8 * (256 + 2*5120) fp32 = 328 KB per GPU, held from library load to process exit); inputs and outputs are still the
 * caller's.  Limit: B * (freq_dim + 2*D) <= 83,968 (e.g. B = 8 at freq_dim 256, D 5120); anything larger returns < 0 with a
 * message that names the limit — use vc_time_embed there.  The call itself performs no cudaMalloc / cudaFree and calls
 * no library that allocates.  The first call on each device looks up the array's device address (with that device
 * current); make that first call outside any stream capture.  All calls on a device share the one array: it
 * keeps whatever the last call wrote, which is harmless because each call overwrites every value before reading it (nothing
 * needs zeroing), but two calls whose kernels can overlap on the GPU (different streams, no ordering between them) race
 * on it and corrupt each other — order them.  Same arithmetic and launches as vc_time_embed: outputs are bit-identical. */
int vc_time_embed_static(const double* t, int B, int freq_dim, int D, const void* w0, const void* b0, const void* w2,
                         const void* b2, const void* wp, const void* bp, void* e_out, void* e0_out, void* stream);

/* a4/a6 — `self.modulation + e` of every block in one launch (wan_transformer3d.py:588 per block, Head :640-643):
 *   out[i, b, :]  = bf16(block_modulations[i][0:6D] + e0[b, 0:6D])      i < n_blocks   (out: [n_blocks, B, 6D])
 *   out_head[b,:] = bf16(head_modulation[0:2D] + (e[b] | e[b]))          when head_modulation != NULL  ([B, 2D])
 * block_modulations: DEVICE array of n_blocks device pointers to the blocks' [1, 6, D] parameters. */
int vc_modulation_table(const void* const* block_modulations, int n_blocks, const void* e0, int B, int D, void* out,
                        const void* head_modulation, const void* e, void* out_head, void* stream);

/* K6/K7 — o = softmax(q k^T * softmax_scale) v, non-causal, all Lk keys attended (no mask other than the tail of
 * the last 128-key tile).  Replaces videox_fun attention() at wan_transformer3d.py:394-399 (self) and :425-430
 * (cross).  head_dim must be 128 and contiguous; *_strides = {batch, token, head} strides in elements (so q/k/v
 * may be column slices of one fused QKV buffer).  variant selects the kernel schedule (same arithmetic):
 *   0: P staged through shared memory (SS MMA);  1: P written over S in tensor memory (TS MMA);
 *   2: one S buffer shared by both Q tiles, P in its own TMEM columns (S_j+1 overlaps softmax_j);
 *   3: 2 + split PV issue + packed fp32x2 softmax;  4 / 5: 3 + 16 / 32 of every 128 exp2 evaluated by a
 *   degree-3 polynomial on the FMA pipe instead of the MUFU (rel. error < 1e-4, below P's bf16 rounding).
 *   4 is the fastest on B200 under the power cap and is what the drop-in module uses for self-attention.
 *   6: short key sequences (cross-attention to the 512 text tokens): one 128-row Q tile per CTA, 256 TMEM columns,
 *   96 KB smem, so two CTAs share an SM and each one's prologue/epilogue hides behind the other's MMAs.
 *   7 / 8: 6 with 64-key tiles and S double-buffered inside the same 256 TMEM columns (S_j+1 is computed during
 *   softmax_j); 8 moves a quarter of the exp2 to the FMA pipe.  9: the pipeline of 7 in a persistent CTA (2 per SM) that
 *   walks Q tiles: barriers / TMEM set up once, next Q prefetched, O drained under the next tile's S MMAs. */
int vc_flash_attn_fwd(const void* q, const void* k, const void* v, void* o, int B, int H, int Lq, int Lk,
                      int head_dim, const long long* q_strides, const long long* k_strides,
                      const long long* v_strides, const long long* o_strides, float softmax_scale, int variant,
                      void* stream);

/* C1 — Ulysses sequence parallelism fused into the producer kernels (replaces the two all-to-alls that the
 * reference gets from xfuser's SeqAllToAll4D inside videox_fun.dist.usp_attn_forward, bound at
 * wan_transformer3d.py:906-921; source not in the reference tree).  The peer pointers are device addresses of the
 * other GPUs' buffers mapped into this process (CUDA peer / symmetric memory); the kernels store straight into them
 * over NVLink.  The caller orders the kernels across GPUs with a stream barrier (see versecrafter_b200/sp.py).
 *
 * vc_rmsnorm_rope_qkv_scatter: rows [q(D) | k(D) | v(D)] (row stride ld) of THIS rank's tokens -> q,k normalised and
 *   rotated exactly as vc_rmsnorm_rope_qk, v copied, and head h of global token t of sample b stored at
 *   peer_recv[h / heads_per_peer] + ((((which*n_batch + b)*l_total + t)*heads_per_peer + h % heads_per_peer)*head_dim,
 *   which = 0,1,2 for q,k,v; n_batch = rows / rows_per_batch; t = pos_offset + row % rows_per_batch.  which_mask
 *   (bit 0 q, bit 1 k, bit 2 v; 7 = all) selects the matrices of this launch, so the caller can scatter q while the
 *   k GEMM still runs (versecrafter_b200/sp.py); bit 3 (8): stores issued by the TMA engine (one cp.async.bulk per row,
 *   matrix and peer) instead of per-thread 16-byte stores; bits 8-15: cap on resident CTAs per SM (0 = default).
 * vc_gemm_bf16_scatter: Y = A . W^T + bias (as vc_gemm_bf16 with VC_EPI_BIAS) whose epilogue stores head h of token row
 *   r at peer_recv[h / heads_per_peer] + (((b*l_total + t)*heads_per_peer + h % heads_per_peer)*head_dim, b = r /
 *   rows_per_batch, t = pos_offset + r % rows_per_batch: the v Linear (wan_transformer3d.py:383-389) and its Ulysses
 *   all-to-all in one kernel.  peer_recv[i] points at the matrix' own block of peer i's buffer; head_dim = 128;
 *   sched = 0 or one VC_GEMM_SCHED_* bit.
 * vc_flash_attn_fwd_scatter: vc_flash_attn_fwd on the local heads over all tokens, with query row t stored at row
 *   t % rows_per_peer of peer_o[t / rows_per_peer] (strides o_strides; the caller bakes its head offset into the
 *   pointers).  variant >= 2. */
int vc_rmsnorm_rope_qkv_scatter(const void* qkv, long long ld, long long rows, long long rows_per_batch,
                                long long pos_offset, int D, const void* gain_q, const void* gain_k,
                                const void* rope_table, int grid_f, int grid_h, int grid_w, int nf, int nh, int nw,
                                float eps, void* const* peer_recv, int n_peers, int heads_per_peer, long long l_total,
                                int which_mask, void* stream);
/* Stream-ordered barrier of the sequence-parallel group over peer-mapped flag arrays (one u32 per peer, zero-initialised):
 * publishes `epoch` (1-based call count, equal on all ranks) into slot [rank] of every peer's array and waits until every
 * peer's epoch has arrived in this rank's array.  Work enqueued after it starts only after the peers' kernels enqueued
 * before their matching call have completed.  A peer that does not arrive within timeout_ms (0 = 5000) sets
 * *status = 0x80000000 | peer instead of hanging the GPU. */
int vc_peer_barrier(void* const* flags, int rank, int n_peers, unsigned int epoch, int timeout_ms, void* status,
                    void* stream);
/* Split form of vc_peer_barrier (same flag arrays, same epoch counter): vc_peer_signal publishes `epoch` to every peer once
 * everything enqueued before it on the stream has completed; vc_peer_wait blocks the stream until every peer has published
 * an epoch >= `epoch`.  Several signals may be outstanding; the host puts independent work between a signal and its wait
 * (the interleaved backbone / GeoAdapter schedule under sequence parallelism) so that a late peer costs nothing.
 * Replaces, like vc_peer_barrier, the synchronisation inside xfuser's all_to_all (wan_transformer3d.py:906-921). */
int vc_peer_signal(void* const* flags, int rank, int n_peers, unsigned int epoch, void* stream);
int vc_peer_wait(void* const* flags, int rank, int n_peers, unsigned int epoch, int timeout_ms, void* status, void* stream);
int vc_gemm_bf16_scatter(const void* A, long long lda, const void* W, long long ldw, long long M, long long N,
                         long long K, const void* bias, long long rows_per_batch, long long pos_offset,
                         void* const* peer_recv, int n_peers, int heads_per_peer, int head_dim, long long l_total,
                         int sched, void* stream);
int vc_flash_attn_fwd_scatter(const void* q, const void* k, const void* v, int B, int H, int Lq, int Lk, int head_dim,
                              const long long* q_strides, const long long* k_strides, const long long* v_strides,
                              void* const* peer_o, int n_peers, int rows_per_peer, const long long* o_strides,
                              float softmax_scale, int variant, void* stream);

/* K1/K2 front end — one sample [C, F, H, W] -> patch matrix A[L, 4C] (row stride ld_out), K ordered (c, ph, pw),
 * token n = (f*H/2 + h)*W/2 + w; Conv3d(kernel = stride = (1,2,2)) == vc_gemm_bf16(A, weight.view(D, 4C)) + bias
 * (wan_transformer3d.py:758-759, wan_transformer3d_versecrafter.py:199-201, :262-263, :340-343). */
int vc_patchify(const void* in, void* out, int C, int F, int H, int W, long long ld_out, void* stream);

/* K1/K2 — the patch-embedding Conv3d(kernel = stride = (1,2,2)) of ONE sample x[C, F, H, W] as a TMA-fed implicit
 * GEMM: out[n - tok_lo, :] = sum_{c,ph,pw} x[c, f, 2h+ph, 2w+pw] * weight[:, (c*2 + ph)*2 + pw] + bias for the tokens
 * n = (f*H/2 + h)*W/2 + w in [tok_lo, tok_hi) (a rank's token range under sequence parallelism).  weight is the
 * Conv3d weight [D, C, 1, 2, 2] viewed [D, 4C] (row stride ldw), out rows are ldo elements apart.  The input is read
 * by 4-D TMA boxes of whole patch rows and re-laid out in shared memory as the swizzled K-major MMA operand; no
 * im2col matrix is written to memory.  Replaces patch_embedding (wan_transformer3d.py:758-759, used
 * wan_transformer3d_versecrafter.py:340-343) and geoada_patch_embedding (wan_transformer3d_versecrafter.py:199-201,
 * :262-267) including flatten(2).transpose(1,2).  Shape support (TMA stride/box limits): C % 16 == 0, C <= 128,
 * H even, W % 8 == 0, W <= 256, D % 8 == 0 — vc_patch_embed_supported returns 1/0; other shapes go through
 * vc_patchify + vc_gemm_bf16. */
int vc_patch_embed_supported(int C, int F, int H, int W, int D);
int vc_patch_embed(const void* x, int C, int F, int H, int W, const void* weight, long long ldw, const void* bias,
                   int D, void* out, long long ldo, long long tok_lo, long long tok_hi, void* stream);

/* a16 back end — y [L, 4*Co] (row stride ldy) -> out [Co, F, 2*gh, 2*gw]  (unpatchify, wan_transformer3d.py:1127-1150) */
int vc_unpatchify(const void* y, long long ldy, void* out, int Co, int gf, int gh, int gw, void* stream);

/* N3 — GeoAdapter input assembly (pipeline_wan_versecrafter.py:397-489) in one pass:
 *   out[s*c_lat + c, f, h, w]            = latents[s][c, f, h, w]                 (torch.cat of the control-video latents)
 *   out[n_src*c_lat + i*8 + j, f, h, w]  = mask[src_frame[f - n_ref], 8h + i, 8w + j]   for f >= n_ref, 0 before
 * i.e. geoada_encode_masks' view/permute/reshape to [64, T, Hl, Wl] followed by F.interpolate(mode='nearest-exact') to
 * Fl - n_ref frames (src_frame = that resampling's source index per output frame, computed by the caller with the
 * same torch routine) and the zero frames prepended for n_ref reference images, then geoada_latent's channel concat.
 * latents: n_src device pointers (host array) to bf16 [c_lat, Fl, Hl, Wl]; mask: bf16 [T, 8*Hl, 8*Wl] (channel 0 of
 * the mask video); out: bf16 [n_src*c_lat + 64, Fl, Hl, Wl]. */
int vc_control_latents(const void* const* latents, int n_src, int c_lat, const void* mask, int T, int Fl, int Hl,
                       int Wl, int n_ref, const int* src_frame, void* out, void* stream);

/* N1 — classifier-free-guidance combine + flow-matching Euler update, one pass over the latents
 * (pipeline_wan_versecrafter.py:903-909; FlowMatchEulerDiscreteScheduler.step of diffusers, not in the reference tree):
 *   guided = bf16(u + bf16(guidance * bf16(c - u)))      u / c = pred_uncond / pred_text (the two halves of the batch)
 *   latents_out = bf16(float(latents) + bf16(dt * guided)),   dt = sigma_next - sigma (the caller passes it already
 *   rounded to bf16: torch's type promotion does that to the 0-dim fp32 dt in `dt * model_output`)
 * pred_text == NULL: no CFG, guided = pred_uncond.  guided_out may be NULL.  n % 8 == 0 elements, bf16. */
int vc_cfg_euler_step(const void* pred_uncond, const void* pred_text, const void* latents, void* latents_out,
                      void* guided_out, long long n, float guidance, float dt, void* stream);

/* N1 — CFG combine + one step of the flow-matching UniPC multistep scheduler (predict-x0, B(h)=expm1(h), solver order
 * <= 2: corrector on the incoming sample, then predictor), one pass over the latents.  Replaces
 * `self.scheduler.step(noise_pred, t, latents)` for the reference's default sampler "Flow_Unipc"
 * (/root/reference/inference/versecrafter_inference.py:135,:254; pipeline_wan_versecrafter.py:903-909; the class is
 * videox_fun.utils.fm_solvers_unipc.FlowUniPCMultistepScheduler, un-vendored).
 *   latents (bf16, n elements) = the sample the model was evaluated on; last/m0/m1 (fp32) = the previous step's
 *   corrected sample and the two most recent x0 predictions (NULL where the order does not need them);
 *   x0_out / last_out (fp32) receive this step's x0 prediction and corrected sample (the caller rotates the history);
 *   coef = 13 host floats {sigma, corr_order, c_x, c_m0, c_bh, c_rho0, c_rk, c_rhot, pred_order, p_x, p_m0, p_bh, p_rk}
 *   (formula in csrc/loop_glue.cu; versecrafter_b200/schedulers.py computes them).  n % 4 == 0. */
int vc_cfg_unipc_step(const void* pred_uncond, const void* pred_text, const void* latents, const void* last,
                      const void* m0, const void* m1, void* latents_out, void* x0_out, void* last_out, long long n,
                      float guidance, const float* coef, void* stream);

/* N4 — per-pixel control-map kernels of /root/reference/inference/rendering_4D_control_maps.py (the parts that are
 * plain tensor arithmetic there; the pytorch3d point/mesh rasterisers are not built).  Images are uint8 [.., H, W, 3],
 * depths fp32 [.., H, W], masks uint8 0/1 (torch.bool storage); n_pix = frames*H*W.
 *
 * vc_composite_by_depth: composite_by_depth (:438-455) + merge_bg_and_fg_mask (:736-763):
 *   take_fg = fg_mask & (bg_depth <= 0 | (fg_depth > 0 & fg_depth < bg_depth - 1e-6));  out_rgb / out_depth = fg where
 *   take_fg else bg;  out_mask_rgb = 255 * (take_fg ? fg_mask : !bg_mask) on 3 channels.  Any output may be NULL.
 * vc_depth_to_gray: visualize_depth_as_grayscale (:487-539) for a given range: disp = depth > 0 ? 1/depth : 0; if
 *   normalise: (disp - min_disp) / denom with min_disp = 1/max_depth, denom = 1/min_depth - 1/max_depth + 1e-8 (computed
 *   by the caller in double, as the reference's python floats); gray = uint8(clamp(.,0,1)*255) on 3 channels.
 * vc_gaussian_projection: one frame of project_3d_gaussians_to_2d (:573-694, density :765-883) + the blend over the
 *   background (:697-733) + main()'s masked projection frame: params[k] = {mean_u, mean_v, inv00, inv01, inv10, inv11,
 *   coeff, r, g, b} (HOST array, Gaussians sorted far -> near, <= 32); density normalised by its image maximum
 *   (max_scratch: >= 32 uint32 of device memory), thresholded to alpha, composited; outputs (each may be NULL):
 *   out_rgb uint8, out_alpha fp32, out_blend uint8 (needs bg_rgb), out_masked uint8 = uint8(rgb/255 * alpha * 255). */
int vc_composite_by_depth(const void* bg_rgb, const void* bg_depth, const void* bg_mask, const void* fg_rgb,
                          const void* fg_depth, const void* fg_mask, void* out_rgb, void* out_depth, void* out_mask_rgb,
                          long long n_pix, void* stream);
int vc_depth_to_gray(const void* depth, void* out_rgb, long long n_pix, float min_disp, float denom, int normalise,
                     void* stream);
int vc_gaussian_projection(int n_gauss, const float* params, int W, int H, float threshold, const void* bg_rgb,
                           void* max_scratch, void* out_rgb, void* out_alpha, void* out_blend, void* out_masked,
                           void* stream);

/* N4 — the two rasterisers the reference delegates to pytorch3d (rendering_4D_control_maps.py:150-338), restated from
 * pytorch3d's published naive algorithms (pytorch3d itself is absent: parity unpinned, oracle/raster_restate.py).
 * Camera: K = 3x3 pixel intrinsics, T = 4x4 OpenCV world-to-camera (both HOST, row-major fp32), converted as
 * _build_cam_from_extrinsics (:338-396) + PerspectiveCameras(in_ndc=False) do.  One call renders ONE frame.
 *
 * vc_render_points: render_point_cloud_pytorch3d_batch (:243-336): points fp32 [n,3] (world), colors uint8 [n,3];
 *   per pixel the 8 nearest points (view z >= 0, NDC distance^2 < radius^2) are alpha-composited front to back with
 *   weights 1 - d^2/r^2, empty pixels get background[3]; out_rgb uint8 [H,W,3], out_depth fp32 [H,W] (nearest z, 0 if
 *   empty), out_mask uint8 [H,W].  Scratch: proj_scratch >= n*3 fp32, sel_scratch >= 8*H*W uint64.
 * vc_render_mesh_depth: depth/mask part of render_meshes_pytorch3d_batch (:150-241) (blur_radius 0, 1 face per pixel,
 *   perspective-correct barycentrics): verts fp32 [n_verts,3] (world), faces int32 [n_faces,3]; out_depth fp32 [H,W],
 *   out_mask uint8 [H,W], out_pix_to_face int32 [H,W] (may be NULL).  Scratch: vproj_scratch >= n_verts*3 fp32,
 *   zkey_scratch >= H*W uint64. */
int vc_render_points(const void* points, const void* colors, long long n_points, const float* K, const float* T, int W,
                     int H, float radius, const float* background, void* proj_scratch, void* sel_scratch, void* out_rgb,
                     void* out_depth, void* out_mask, void* stream);
int vc_render_mesh_depth(const void* verts, int n_verts, const void* faces, int n_faces, const float* K, const float* T,
                         int W, int H, void* vproj_scratch, void* zkey_scratch, void* out_depth, void* out_mask,
                         void* out_pix_to_face, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VC_B200_H */
