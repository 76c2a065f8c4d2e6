"""GPU parity of each sm_100a kernel, called through the C ABI (ctypes), against the oracle's functions
(oracle/restate.py) on the same seeded inputs.  Tolerances: bf16 outputs, so "equal up to one bf16 rounding":
relative L2 <= 4e-3 for the row kernels and GEMM epilogues (measured ~1e-3 when a rounding tie flips), <= 1e-2
for attention (P is rounded to bf16 before the PV product, as in flash-attn)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import restate  # noqa: E402


def relerr(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


@pytest.fixture(scope="module")
def ops():
    from versecrafter_b200 import _lib, ops
    lib = _lib.load()
    assert lib.vc_device_ok() == 0, lib.vc_last_error()
    return ops


def _rand(shape, seed, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(shape, generator=g, device="cuda") * scale).to(torch.bfloat16)


@pytest.mark.parametrize("D", [128, 256, 512, 1536, 2048, 5120])
@pytest.mark.parametrize("rows_per_batch", [1, 77, 333])
def test_ln_modulate(ops, D, rows_per_batch):
    B = 2
    x = _rand((B, rows_per_batch, D), 1, 2.0) + 0.5
    m = _rand((B, 6, D), 2, 0.3)
    out = ops.ln_modulate(x, m[:, 0], m[:, 1], rows_per_batch)
    ref = restate.layer_norm(x) * (1 + m[:, 1:2]) + m[:, 0:1]
    assert out.shape == x.shape
    assert relerr(out, ref) < 4e-3
    assert (out.float() - ref.float()).abs().max() <= 0.07 * ref.float().abs().max()


@pytest.mark.parametrize("D", [1032, 3072, 4096, 6144, 8192])
def test_ln_modulate_wide_rows(ops, D):
    """Block-per-row kernel (D > 1024) at every vectors-per-thread instantiation, incl. a ragged last vector group,
    and more rows than resident CTAs so the grid-stride loop (and its shared-memory reuse) is exercised."""
    B, rpb = 2, 2500
    x = _rand((B, rpb, D), 21, 2.0) + 0.5
    m = _rand((B, 6, D), 22, 0.3)
    out = ops.ln_modulate(x, m[:, 3], m[:, 4], rpb)
    ref = restate.layer_norm(x) * (1 + m[:, 4:5]) + m[:, 3:4]
    assert relerr(out, ref) < 4e-3
    gamma = (1 + 0.1 * torch.randn(D, device="cuda")).to(torch.bfloat16)
    beta = (0.02 * torch.randn(D, device="cuda")).to(torch.bfloat16)
    assert relerr(ops.ln_affine(x, gamma, beta), restate.layer_norm(x, gamma, beta)) < 4e-3


@pytest.mark.parametrize("D", [128, 1536, 5120])
def test_ln_affine(ops, D):
    x = _rand((3, 50, D), 3, 1.5)
    gamma = (1 + 0.1 * torch.randn(D, device="cuda")).to(torch.bfloat16)
    beta = (0.02 * torch.randn(D, device="cuda")).to(torch.bfloat16)
    out = ops.ln_affine(x, gamma, beta)
    ref = restate.layer_norm(x, gamma, beta)
    assert relerr(out, ref) < 4e-3


@pytest.mark.parametrize("D,H", [(128, 1), (256, 2), (1536, 12), (5120, 40)])
@pytest.mark.parametrize("grid,pos_offset,rows_per_batch", [((3, 4, 6), 0, 72), ((3, 4, 6), 36, 36), ((2, 3, 5), 0, 37)])
def test_rmsnorm_rope(ops, D, H, grid, pos_offset, rows_per_batch):
    """q,k normalised over the full D then rotated per head; v untouched; pad rows (pos >= F*H*W) not rotated;
    pos_offset emulates a sequence-parallel rank that owns the second half of the tokens."""
    B = 2
    d = D // H
    qkv = _rand((B * rows_per_batch, 3 * D), 4)
    gq = (1 + 0.1 * torch.randn(D, device="cuda")).to(torch.bfloat16)
    gk = (1 + 0.1 * torch.randn(D, device="cuda")).to(torch.bfloat16)
    ref_in = qkv.clone()
    rope = ops.rope_table(d, grid, "cuda")
    ops.rmsnorm_rope_(qkv, D, 2, gq, gk, rows_per_batch, rope=rope, grid=grid, pos_offset=pos_offset)
    ang = restate.rope_angles(d, grid, device="cuda")
    L = grid[0] * grid[1] * grid[2]
    for i, g in enumerate((gq, gk)):
        u = restate.rms_norm(ref_in[:, i * D:(i + 1) * D], g).view(B, rows_per_batch, H, d)
        # oracle rotates rows [0, n) of a sequence that starts at position 0: build the full sequence, then slice
        full = torch.zeros(B, max(L, pos_offset + rows_per_batch), H, d, dtype=u.dtype, device="cuda")
        full[:, pos_offset:pos_offset + rows_per_batch] = u
        want = restate.rope_rotate(full, ang)[:, pos_offset:pos_offset + rows_per_batch]
        got = qkv[:, i * D:(i + 1) * D].view(B, rows_per_batch, H, d)
        assert relerr(got, want) < 4e-3, i
    assert torch.equal(qkv[:, 2 * D:], ref_in[:, 2 * D:])


@pytest.mark.parametrize("D,H", [(3072, 24), (4096, 32), (8192, 64)])
def test_rmsnorm_rope_wide_rows(ops, D, H):
    """Block-per-row kernel with the grid-stride loop active (rows > resident CTAs), q and k, SP position offset."""
    B, grid, pos_offset = 2, (5, 20, 30), 1500
    rpb, d = 1500, D // H
    qkv = _rand((B * rpb, 3 * D), 23)
    gq = (1 + 0.1 * torch.randn(D, device="cuda")).to(torch.bfloat16)
    gk = (1 + 0.1 * torch.randn(D, device="cuda")).to(torch.bfloat16)
    ref_in = qkv.clone()
    rope = ops.rope_table(d, grid, "cuda")
    ops.rmsnorm_rope_(qkv, D, 2, gq, gk, rpb, rope=rope, grid=grid, pos_offset=pos_offset)
    ang = restate.rope_angles(d, grid, device="cuda")
    for i, g in enumerate((gq, gk)):
        u = restate.rms_norm(ref_in[:, i * D:(i + 1) * D], g).view(B, rpb, H, d)
        full = torch.zeros(B, pos_offset + rpb, H, d, dtype=u.dtype, device="cuda")
        full[:, pos_offset:] = u
        want = restate.rope_rotate(full, ang)[:, pos_offset:]
        assert relerr(qkv[:, i * D:(i + 1) * D].view(B, rpb, H, d), want) < 4e-3, i
    assert torch.equal(qkv[:, 2 * D:], ref_in[:, 2 * D:])


def test_rmsnorm_only(ops):
    D = 1536
    buf = _rand((100, 2 * D), 5)
    g = (1 + 0.1 * torch.randn(D, device="cuda")).to(torch.bfloat16)
    ref = restate.rms_norm(buf[:, :D].clone(), g)
    keep = buf[:, D:].clone()
    ops.rmsnorm_rope_(buf, D, 1, g, None, 100)
    assert relerr(buf[:, :D], ref) < 4e-3
    assert torch.equal(buf[:, D:], keep)


def _gemm_ref(a, w, bias, epi, gate=None, rpb=1, res=None, scale=1.0):
    bf = lambda t: t.to(torch.bfloat16).float()
    v = a.float() @ w.float().t()
    if epi & 1:
        v = v + bias.float()
    if epi & ~1:
        v = bf(v)
    if epi & 2:
        v = bf(torch.nn.functional.gelu(v, approximate="tanh"))
    if epi & 4:
        rows = torch.arange(a.shape[0], device=a.device) // rpb
        v = bf(v * gate.float()[rows])
    if epi & 16:
        v = bf(v * scale)
    if epi & 8:
        v = v + res.float()
    return v.to(torch.bfloat16)


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (128, 256, 256), (300, 512, 512), (1000, 768, 1536), (77, 64, 128),
                                   (4095, 1536, 512), (2, 512, 256)])
def test_gemm_bias(ops, M, N, K):
    a, w, b = _rand((M, K), 6), _rand((N, K), 7, 1 / math.sqrt(K)), _rand((N,), 8, 0.1)
    out = ops.gemm(a, w, b)
    assert relerr(out, _gemm_ref(a, w, b, 1)) < 4e-3
    out0 = ops.gemm(a, w, None)
    assert relerr(out0, _gemm_ref(a, w, None, 0)) < 4e-3


@pytest.mark.parametrize("epi", [3, 9, 13, 25])
def test_gemm_epilogues(ops, epi):
    M, N, K, rpb = 650, 768, 512, 325  # a 128-row tile straddles the batch boundary at row 325
    a, w, b = _rand((M, K), 9), _rand((N, K), 10, 1 / math.sqrt(K)), _rand((N,), 11, 0.1)
    gate = _rand((2, 6, N), 12, 0.5)[:, 2]  # strided [B, N] view like m[:, 2]
    res = _rand((M, N), 13)
    out = ops.gemm(a, w, b, epilogue=epi, gate=gate if epi & 4 else None, rows_per_batch=rpb,
                   residual=res if epi & 8 else None, scale=0.6)
    ref = _gemm_ref(a, w, b, epi, gate, rpb, res, 0.6)
    assert relerr(out, ref) < 4e-3


@pytest.mark.parametrize("M,N,K", [(650, 768, 512), (129, 256, 64), (1950, 1536, 2048), (4096, 5120, 512)])
@pytest.mark.parametrize("epi", [1, 3, 13])
def test_gemm_multicast_schedule_is_bit_identical(ops, M, N, K, epi):
    """The cluster-of-2 TMA-multicast schedule (odd and even M-tile counts, N tails) gives the plain kernel's bits."""
    a, w, b = _rand((M, K), 31), _rand((N, K), 32, 1 / math.sqrt(K)), _rand((N,), 33, 0.1)
    gate, res = _rand((2, N), 34, 0.5), _rand((M, N), 35)
    kw = dict(gate=gate if epi & 4 else None, rows_per_batch=(M + 1) // 2, residual=res if epi & 8 else None)
    plain = ops.gemm(a, w, b, epilogue=epi | ops.SCHED_SINGLE, **kw)
    mc = ops.gemm(a, w, b, epilogue=epi | ops.SCHED_MULTICAST, **kw)
    assert torch.equal(plain, mc)
    assert relerr(mc, _gemm_ref(a, w, b, epi, gate, (M + 1) // 2, res, 1.0)) < 4e-3


@pytest.mark.parametrize("M,N,K", [(650, 768, 512), (129, 256, 64), (1950, 1536, 2048), (4096, 5120, 512), (256, 64, 128),
                                   (8190, 1536, 1536), (300, 520, 72)])
@pytest.mark.parametrize("epi", [0, 1, 3, 9, 13, 25])
def test_gemm_cta_pair_schedule_is_bit_identical(ops, M, N, K, epi):
    """cta_group::2: one tcgen05.mma of M = 256 per CTA pair, each CTA holding its 128 A rows and half of the W tile (odd
    and even M-tile counts, N and K tails, every epilogue) — same bits as the one-CTA-per-tile kernel."""
    a, w, b = _rand((M, K), 36), _rand((N, K), 37, 1 / math.sqrt(K)), _rand((N,), 38, 0.1)
    gate, res = _rand((2, N), 39, 0.5), _rand((M, N), 40)
    kw = dict(gate=gate if epi & 4 else None, rows_per_batch=(M + 1) // 2, residual=res if epi & 8 else None, scale=0.6)
    bias = b if epi & 1 else None
    plain = ops.gemm(a, w, bias, epilogue=epi | ops.SCHED_SINGLE, **kw)
    pair = ops.gemm(a, w, bias, epilogue=epi | ops.SCHED_PAIR, **kw)
    assert torch.equal(plain, pair)
    assert relerr(pair, _gemm_ref(a, w, bias, epi, gate, (M + 1) // 2, res, 0.6)) < 4e-3


def test_gemm_cta_pair_many_waves(ops):
    """Persistent pair schedule over many waves with both accumulator stages and all six smem stages cycling."""
    M, N, K = 148 * 128 * 2 + 77, 1024, 1024
    a, w, b = _rand((M, K), 41), _rand((N, K), 42, 1 / math.sqrt(K)), _rand((N,), 43, 0.1)
    assert torch.equal(ops.gemm(a, w, b, epilogue=1 | ops.SCHED_SINGLE), ops.gemm(a, w, b, epilogue=1 | ops.SCHED_PAIR))


def test_gemm_inplace_residual_and_strided_views(ops):
    """x += scale * (c W^T + b) written over x (adapter accumulate) and A given as a column slice of a wider buffer."""
    M, N, K = 500, 512, 256
    wide = _rand((M, 3 * K), 14)
    a = wide[:, K:2 * K]
    w, b = _rand((N, K), 15, 1 / math.sqrt(K)), _rand((N,), 16, 0.1)
    x = _rand((M, N), 17)
    ref = _gemm_ref(a, w, b, 25, res=x.clone(), scale=1.0)
    ops.gemm(a, w, b, epilogue=25, residual=x, scale=1.0, out=x)
    assert relerr(x, ref) < 4e-3


def test_gemm_large_k_many_tiles(ops):
    M, N, K = 2048, 2048, 8960
    a, w, b = _rand((M, K), 18), _rand((N, K), 19, 1 / math.sqrt(K)), _rand((N,), 20, 0.1)
    assert relerr(ops.gemm(a, w, b), _gemm_ref(a, w, b, 1)) < 4e-3


ATTN_SHAPES = [(1, 1, 128, 128), (2, 2, 256, 256), (2, 2, 333, 333), (1, 3, 1950, 1950), (2, 2, 333, 512),
               (1, 2, 4095, 4095), (1, 1, 100, 37)]


def _attn_case(ops, B, H, Lq, Lk, variant, fused):
    d = 128
    if fused and Lq == Lk:  # q/k/v as column slices of one [B*L, 3*H*d] buffer, like the fused QKV GEMM output
        buf = _rand((B * Lq, 3 * H * d), 21)
        q, k, v = (buf[:, i * H * d:(i + 1) * H * d].view(B, Lq, H, d) for i in range(3))
    else:
        q, k, v = _rand((B, Lq, H, d), 22), _rand((B, Lk, H, d), 23), _rand((B, Lk, H, d), 24)
    out = ops.flash_attn(q, k, v, variant=variant)
    ref = restate.sdpa(q, k, v)
    torch.cuda.synchronize()
    return relerr(out, ref)


@pytest.mark.parametrize("B,H,Lq,Lk", ATTN_SHAPES)
@pytest.mark.parametrize("fused", [False, True])
def test_flash_attn_smem_p(ops, B, H, Lq, Lk, fused):
    assert _attn_case(ops, B, H, Lq, Lk, 0, fused) < 1e-2


@pytest.mark.parametrize("B,H,Lq,Lk", ATTN_SHAPES)
def test_flash_attn_tmem_p(ops, B, H, Lq, Lk):
    assert _attn_case(ops, B, H, Lq, Lk, 1, True) < 1e-2


@pytest.mark.parametrize("B,H,Lq,Lk", ATTN_SHAPES + [(1, 2, 1000, 2600), (2, 1, 640, 128)])
@pytest.mark.parametrize("fused", [False, True])
@pytest.mark.parametrize("variant", [2, 3, 4, 5])
def test_flash_attn_shared_s(ops, B, H, Lq, Lk, fused, variant):
    """variants 2-5: one S buffer shared by both Q tiles, P in its own TMEM columns (odd/even tile counts, tails);
    3 = split PV issue + packed fp32x2 softmax, 4 / 5 = 16 / 32 of every 128 exp2 on the FMA pipe (polynomial)."""
    assert _attn_case(ops, B, H, Lq, Lk, variant, fused) < 1e-2


@pytest.mark.parametrize("B,H,Lq,Lk", ATTN_SHAPES + [(2, 3, 1000, 512), (1, 2, 130, 1024), (2, 1, 640, 128), (1, 1, 7, 300)])
@pytest.mark.parametrize("fused", [False, True])
def test_flash_attn_short_keys_x1(ops, B, H, Lq, Lk, fused):
    """variant 6: one 128-row Q tile per CTA, two CTAs per SM (the cross-attention kernel); any Lq / Lk tails."""
    assert _attn_case(ops, B, H, Lq, Lk, 6, fused) < 1e-2


@pytest.mark.parametrize("B,H,Lq,Lk", ATTN_SHAPES + [(2, 3, 1000, 512), (1, 2, 130, 1024), (2, 1, 640, 128), (1, 1, 7, 300),
                                                     (1, 2, 200, 64), (1, 1, 129, 65), (2, 2, 256, 1)])
@pytest.mark.parametrize("fused", [False, True])
@pytest.mark.parametrize("variant", [7, 8, 9])
def test_flash_attn_short_keys_x2(ops, B, H, Lq, Lk, fused, variant):
    """variants 7 / 8: 64-key tiles with a double-buffered S (S_{j+1} computed during softmax_j), one 128-row Q tile per
    CTA, two CTAs per SM; 8 = a quarter of the exponentials on the FMA pipe; 9 = the same pipeline in a persistent CTA
    that walks Q tiles.  Odd / even / single tile counts, tails."""
    assert _attn_case(ops, B, H, Lq, Lk, variant, fused) < 1e-2


@pytest.mark.parametrize("Lk", [512, 192, 65])
def test_flash_attn_persistent_x3_many_tiles_per_cta(ops, Lk):
    """variant 9 with far more Q tiles than resident CTAs (2 per SM), so every CTA walks several tiles: odd and even key-tile
    counts flip the ring parity between consecutive tiles; bit-identical to variant 7 (same arithmetic per tile)."""
    B, H, Lq, d = 2, 6, 128 * 61 + 37, 128          # 2 * 6 * 62 = 744 tiles on 296 CTAs
    q, k, v = _rand((B, Lq, H, d), 31), _rand((B, Lk, H, d), 32), _rand((B, Lk, H, d), 33)
    o9 = ops.flash_attn(q, k, v, variant=9)
    o7 = ops.flash_attn(q, k, v, variant=7)
    assert torch.equal(o9, o7)
    assert relerr(o9, restate.sdpa(q, k, v)) < 1e-2


def test_flash_attn_large_scores_rescale(ops):
    """Rows whose running max keeps growing across key tiles exercise the O rescale path."""
    B, H, L, d = 1, 1, 1024, 128
    q = _rand((B, L, H, d), 25, 3.0)
    k = _rand((B, L, H, d), 26, 1.0)
    k = (k.float() * torch.linspace(0.2, 3.0, L, device="cuda")[None, :, None, None]).to(torch.bfloat16)
    v = _rand((B, L, H, d), 27)
    for variant in (0, 1, 2, 3, 4, 5, 6, 7, 8, 9):
        assert relerr(ops.flash_attn(q, k, v, variant=variant), restate.sdpa(q, k, v)) < 1e-2, variant


@pytest.mark.parametrize("P,H,grid", [(2, 16, (2, 8, 24)), (4, 16, (2, 8, 24)), (8, 16, (2, 8, 24)),
                                      (8, 40, (3, 13, 40)),    # c3's head split (5 heads per rank), Lp = 195: ragged
                                      (4, 40, (5, 6, 12))])    # Q tiles and a token scatter that splits inside a tile
def test_ulysses_fused_kernels_emulated_ranks(ops, P, H, grid):
    """The fused Ulysses kernels with the P ranks emulated on ONE GPU (the "peer" buffers are P local tensors): the
    head-scatter inside RMSNorm+RoPE and the token-scatter inside the attention epilogue must reproduce, bit for bit,
    the un-sharded kernels' results sliced the way the all-to-alls would slice them."""
    B, d = 2, 128
    D, L = H * d, grid[0] * grid[1] * grid[2]
    assert L % P == 0
    Lp, Hp = L // P, H // P
    full = _rand((B, L, 3 * D), 40)
    gq = (1 + 0.1 * torch.randn(D, device="cuda")).to(torch.bfloat16)
    gk = (1 + 0.1 * torch.randn(D, device="cuda")).to(torch.bfloat16)
    rope = ops.rope_table(d, grid, "cuda")
    # un-sharded reference through the plain kernels
    ref = full.reshape(B * L, 3 * D).clone()
    ops.rmsnorm_rope_(ref, D, 2, gq, gk, L, rope=rope, grid=grid)
    q_ref, k_ref, v_ref = (ref[:, i * D:(i + 1) * D].view(B, L, H, d) for i in range(3))
    o_ref = ops.flash_attn(q_ref, k_ref, v_ref, variant=4)
    # emulated ranks
    recv = [torch.full((3, B, L, Hp, d), float("nan"), device="cuda", dtype=torch.bfloat16) for _ in range(P)]
    attn = [torch.full((B, Lp, H, d), float("nan"), device="cuda", dtype=torch.bfloat16) for _ in range(P)]
    for r in range(P):
        mine = full[:, r * Lp:(r + 1) * Lp].reshape(B * Lp, 3 * D).contiguous()
        ops.rmsnorm_rope_qkv_scatter(mine, D, gq, gk, Lp, [t.data_ptr() for t in recv], Hp, L, rope, grid,
                                     pos_offset=r * Lp)
    for p in range(P):
        for i, want in enumerate((q_ref, k_ref, v_ref)):
            assert torch.equal(recv[p][i], want[:, :, p * Hp:(p + 1) * Hp]), (p, i)
    for p in range(P):
        ops.flash_attn_scatter(recv[p][0], recv[p][1], recv[p][2], [t.data_ptr() + p * Hp * d * 2 for t in attn], Lp,
                               (Lp * D, D, d), variant=4)
    for r in range(P):
        assert torch.equal(attn[r], o_ref[:, r * Lp:(r + 1) * Lp]), r
    # zero-padded tail tokens (seq_len > F*H'*W' or L % P != 0) are not attended: keys sliced to the true length
    Lk = L - 5
    o_cut = ops.flash_attn(q_ref, k_ref[:, :Lk], v_ref[:, :Lk], variant=4)
    for p in range(P):
        ops.flash_attn_scatter(recv[p][0], recv[p][1][:, :Lk], recv[p][2][:, :Lk],
                               [t.data_ptr() + p * Hp * d * 2 for t in attn], Lp, (Lp * D, D, d), variant=4)
    for r in range(P):
        assert torch.equal(attn[r], o_cut[:, r * Lp:(r + 1) * Lp]), r


@pytest.mark.parametrize("P,H,grid,K", [(2, 4, (2, 8, 24), 256), (8, 40, (3, 13, 40), 512), (4, 8, (5, 6, 12), 320)])
def test_ulysses_overlapped_pieces_emulated_ranks(ops, P, H, grid, K):
    """The pieces of sp.ulysses_qkv_attention_overlapped with the ranks emulated on one GPU: q and k scattered by separate
    masked launches of the RMSNorm+RoPE scatter kernel, v = Linear(h) scattered by the GEMM's own epilogue
    (vc_gemm_bf16_scatter, every schedule) — bit-identical to the one-launch scatter of the fused QKV GEMM's output."""
    B, d = 2, 128
    D, L = H * d, grid[0] * grid[1] * grid[2]
    Lp, Hp = L // P, H // P
    h = _rand((B, L, K), 60)
    w = _rand((3 * D, K), 61, 1 / math.sqrt(K))
    bias = _rand((3 * D,), 62)
    gq = (1 + 0.1 * torch.randn(D, device="cuda")).to(torch.bfloat16)
    gk = (1 + 0.1 * torch.randn(D, device="cuda")).to(torch.bfloat16)
    rope = ops.rope_table(d, grid, "cuda")
    want = [torch.full((3, B, L, Hp, d), float("nan"), device="cuda", dtype=torch.bfloat16) for _ in range(P)]
    for r in range(P):   # the one-kernel path: fused QKV GEMM, then one scatter launch
        qkv = ops.gemm(h[:, r * Lp:(r + 1) * Lp].reshape(B * Lp, K).contiguous(), w, bias)
        ops.rmsnorm_rope_qkv_scatter(qkv, D, gq, gk, Lp, [t.data_ptr() for t in want], Hp, L, rope, grid, pos_offset=r * Lp)
    for sched, bulk, cap in ((0, False, 0), (ops.SCHED_SINGLE, True, 0), (ops.SCHED_PAIR, True, 1), (ops.SCHED_MULTICAST, False, 2)):
        got = [torch.full((3, B, L, Hp, d), float("nan"), device="cuda", dtype=torch.bfloat16) for _ in range(P)]
        ptrs = [t.data_ptr() for t in got]
        v_block = 2 * B * L * Hp * d * 2
        for r in range(P):
            hr = h[:, r * Lp:(r + 1) * Lp].reshape(B * Lp, K).contiguous()
            qkv = torch.full((B * Lp, 3 * D), float("nan"), device="cuda", dtype=torch.bfloat16)
            ops.gemm(hr, w[:D], bias[:D], out=qkv[:, :D])
            ops.rmsnorm_rope_qkv_scatter(qkv, D, gq, gk, Lp, ptrs, Hp, L, rope, grid, pos_offset=r * Lp, which_mask=1, bulk=bulk,
                                         cta_cap=cap)
            ops.gemm(hr, w[D:2 * D], bias[D:2 * D], out=qkv[:, D:2 * D])
            ops.rmsnorm_rope_qkv_scatter(qkv, D, gq, gk, Lp, ptrs, Hp, L, rope, grid, pos_offset=r * Lp, which_mask=2, bulk=bulk,
                                         cta_cap=cap)
            ops.gemm_scatter(hr, w[2 * D:], bias[2 * D:], Lp, [p_ + v_block for p_ in ptrs], Hp, L, pos_offset=r * Lp,
                             sched=sched)
            assert torch.isnan(qkv[:, 2 * D:].float()).all()     # v never exists as a local matrix
        for p_ in range(P):
            for i in range(3):
                assert torch.equal(got[p_][i], want[p_][i]), (sched, p_, i)
    # the one-launch scatter (all three matrices) with bulk stores
    got = [torch.full((3, B, L, Hp, d), float("nan"), device="cuda", dtype=torch.bfloat16) for _ in range(P)]
    for r in range(P):
        qkv = ops.gemm(h[:, r * Lp:(r + 1) * Lp].reshape(B * Lp, K).contiguous(), w, bias)
        ops.rmsnorm_rope_qkv_scatter(qkv, D, gq, gk, Lp, [t.data_ptr() for t in got], Hp, L, rope, grid, pos_offset=r * Lp, bulk=True)
    for p_ in range(P):
        assert torch.equal(got[p_], want[p_]), p_


def test_gemm_rejects_output_over_its_own_input(ops):
    """The persistent kernel re-reads A rows after C tiles of those rows are stored: A aliasing C is an argument error
    (before_proj wrote over its input); disjoint column blocks of one buffer are allowed."""
    from versecrafter_b200._lib import VCError
    a, w = _rand((512, 512), 50), _rand((512, 512), 51, 1 / math.sqrt(512))
    with pytest.raises(VCError, match="overlap"):
        ops.gemm(a, w, None, out=a)
    wide = _rand((512, 1024), 52)
    want = _gemm_ref(wide[:, :512].contiguous(), w, None, 0)
    ops.gemm(wide[:, :512], w, None, out=wide[:, 512:])
    assert relerr(wide[:, 512:], want) < 4e-3


def test_patchify_unpatchify(ops):
    u = _rand((12, 3, 8, 10), 28)
    got = ops.patchify(u)
    assert torch.equal(got, restate.patchify(u))
    y = _rand((3 * 4 * 5, 64), 29)
    assert torch.equal(ops.unpatchify(y, (3, 4, 5)), restate.unpatchify(y, (3, 4, 5)))


@pytest.mark.parametrize("C,Fr,H,W,D", [(128, 3, 16, 24, 512), (16, 2, 12, 104, 256), (128, 2, 6, 160, 128),
                                        (16, 5, 30, 56, 1536), (128, 1, 2, 256, 256)])
def test_patch_embed_implicit_gemm(ops, C, Fr, H, W, D):
    """Conv3d(kernel = stride = (1,2,2)) as the TMA-fed implicit GEMM vs torch's conv3d in fp32 on the same bf16
    values (wan_transformer3d_versecrafter.py:262-263, :340-343), vs the explicit patchify + GEMM route, and on
    token ranges that do not start on a patch-row boundary (a sequence-parallel rank's chunk)."""
    assert ops.patch_embed_supported(C, Fr, H, W, D)
    u = _rand((C, Fr, H, W), 40)
    wt = _rand((D, C, 1, 2, 2), 41, 1.0 / math.sqrt(4 * C))
    b = _rand((D,), 42, 0.1)
    L = Fr * (H // 2) * (W // 2)
    want = torch.nn.functional.conv3d(u.float()[None], wt.float(), b.float(), stride=(1, 2, 2))[0].flatten(1).t()
    out = torch.full((L, D), float("nan"), device="cuda", dtype=torch.bfloat16)
    ops.patch_embed(u, wt.view(D, -1), b, out)
    assert torch.isfinite(out.float()).all()
    assert relerr(out, want) < 4e-3
    explicit = ops.gemm(ops.patchify(u), wt.view(D, -1), b)
    assert relerr(out, explicit) < 1e-3
    for lo, hi in ((0, L), (37 % L, max(L - 19, 37 % L + 1)), (L // 3, L // 3 + 1), (L - 5, L)):
        part = torch.full((hi - lo + 3, D), 7.0, device="cuda", dtype=torch.bfloat16)
        ops.patch_embed(u, wt.view(D, -1), b, part, lo, hi)
        assert torch.equal(part[:hi - lo], out[lo:hi]), (lo, hi)
        assert (part[hi - lo:] == 7.0).all()          # nothing stored outside the requested range


def test_patch_embed_unsupported_shapes_are_refused(ops):
    from versecrafter_b200 import _lib
    assert not ops.patch_embed_supported(16, 5, 30, 52, 512)    # W*2 bytes is not a 16-byte multiple (c1's latent)
    assert not ops.patch_embed_supported(12, 2, 8, 16, 128)     # C not a multiple of 16
    u, wt, b = _rand((16, 5, 30, 52), 43), _rand((512, 64), 44), _rand((512,), 45)
    with pytest.raises(_lib.VCError):
        ops.patch_embed(u, wt, b, torch.empty(5 * 15 * 26, 512, device="cuda", dtype=torch.bfloat16))


def test_errors_are_reported(ops):
    from versecrafter_b200 import _lib
    a, w = _rand((16, 60), 30), _rand((64, 60), 31)
    with pytest.raises(_lib.VCError):
        ops.gemm(a, w, None)  # K % 8 != 0
    with pytest.raises(_lib.VCError):
        ops.gemm(a.cpu(), w.cpu(), None)  # no CPU fallback


@pytest.mark.parametrize("B,D,freq_dim,tvals", [(2, 512, 256, (500.0, 500.0)), (2, 1536, 256, (999.0, 3.0)),
                                                 (1, 5120, 256, (731.0,)), (3, 256, 64, (0.0, 17.5, 1000.0))])
def test_time_embed_island(ops, B, D, freq_dim, tvals):
    """a4: fp64 sinusoid -> time_embedding -> time_projection, fp32 on bf16-valued weights, vs the oracle (pinned to the
    reference's fp32 island).  Only the fp32 summation order differs from cuBLAS: outputs agree to the last bf16 bit
    almost everywhere."""
    sd = {"time_embedding.0.weight": _rand((D, freq_dim), 60, 0.02), "time_embedding.0.bias": _rand((D,), 61, 0.02),
          "time_embedding.2.weight": _rand((D, D), 62, 0.02), "time_embedding.2.bias": _rand((D,), 63, 0.02),
          "time_projection.1.weight": _rand((6 * D, D), 64, 1 / math.sqrt(D)), "time_projection.1.bias": _rand((6 * D,), 65, 0.02)}
    t = torch.tensor(tvals, device="cuda")
    e_ref, e0_ref = restate.time_embed(sd, t, freq_dim, D, torch.bfloat16)
    for tt in (t, t.double(), t.long() if all(float(v).is_integer() for v in tvals) else t):
        e, e0 = ops.time_embed(tt, sd["time_embedding.0.weight"], sd["time_embedding.0.bias"], sd["time_embedding.2.weight"],
                               sd["time_embedding.2.bias"], sd["time_projection.1.weight"], sd["time_projection.1.bias"])
        assert e.shape == (B, D) and e0.shape == (B, 6, D)
        assert relerr(e, e_ref) < 2e-3 and relerr(e0, e0_ref) < 2e-3
        assert float((e0 != e0_ref).float().mean()) < 0.02 and float((e != e_ref).float().mean()) < 0.02


def test_modulation_table_is_the_bf16_add(ops):
    """m[i, b] = modulation_i + e0[b] for all blocks in one launch: bit-identical to torch's bf16 add."""
    B, D, n = 2, 640, 5
    mods = [_rand((1, 6, D), 70 + i, 0.05) for i in range(n)]
    e0, e, head = _rand((B, 6, D), 80, 0.3), _rand((B, D), 81, 0.3), _rand((1, 2, D), 82, 0.05)
    ptrs = torch.tensor([m.data_ptr() for m in mods], dtype=torch.int64).cuda()
    out = torch.empty((n, B, 6, D), device="cuda", dtype=torch.bfloat16)
    out_head = torch.empty((B, 2, D), device="cuda", dtype=torch.bfloat16)
    ops.modulation_table(ptrs, n, e0, out, head, e, out_head)
    for i in range(n):
        assert torch.equal(out[i], mods[i] + e0)
    assert torch.equal(out_head, head + e[:, None, :])
    out2 = torch.zeros_like(out)
    ops.modulation_table(ptrs, n, e0, out2)          # without the head table
    assert torch.equal(out2, out)


@pytest.mark.parametrize("rows,D,rpb", [(7, 1536, 4), (1950, 1536, 975), (5000, 5120, 2500), (3, 5120, 2), (4100, 8192, 4100),
                                        (2221, 2048, 1111), (900, 3072, 450)])
def test_row_ring_kernels_bit_identical_to_register_kernels(ops, rows, D, rpb):
    """The asynchronous row-ring kernels (cp.async.bulk prefetch, round 2) and the register-resident kernels (round 1) run
    the same arithmetic: LayerNorm+modulate, LayerNorm affine and RMSNorm(+RoPE) agree bit for bit — fewer rows than
    CTAs, many rows per CTA, row-strided inputs."""
    wide = _rand((rows, D + 64), 90)
    x = wide[:, :D]                                       # row stride D + 64
    m = _rand((2, 6, D), 91, 0.1)
    a = ops.ln_modulate(x, m[:, 0], m[:, 1], rpb, ring=True)
    b = ops.ln_modulate(x, m[:, 0], m[:, 1], rpb, regs=True)
    assert torch.equal(a, b)
    want = restate.layer_norm(x[:rpb]) * (1 + m[0, 1]) + m[0, 0]
    assert relerr(a[:rpb], want) < 4e-3
    g, bt = _rand((D,), 92, 0.5), _rand((D,), 93, 0.1)
    assert torch.equal(ops.ln_affine(x, g, bt, ring=True), ops.ln_affine(x, g, bt, regs=True))
    assert torch.equal(ops.ln_affine(x, g, bt), ops.ln_affine(x, g, bt, regs=True))      # whatever the default picks
    H = D // 128
    grid = (2, 5, (rpb + 9) // 10)
    rope = ops.rope_table(128, grid, "cuda")
    for n_mats, use_rope in ((2, True), (1, False)):
        q1 = _rand((rows, 3 * D), 94)
        q2 = q1.clone()
        kw = dict(rope=rope, grid=grid) if use_rope else {}
        ops.rmsnorm_rope_(q1, D, n_mats, g, bt, rpb, pos_offset=3, ring=True, **kw)
        ops.rmsnorm_rope_(q2, D, n_mats, g, bt, rpb, pos_offset=3, regs=True, **kw)
        assert torch.equal(q1, q2), (n_mats, use_rope)
