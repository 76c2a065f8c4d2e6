"""Out-of-bounds evidence without compute-sanitizer (closed on this GPU pool: profiles/sanitizer_r02.txt).  Every output
of a kernel family lives in the middle of a larger allocation whose guard bands (64 KB on either side, and — for strided
outputs — the unused tail of every row) hold a sentinel; every input sits between guard bands of NaN.  After the launch
the sentinels must be untouched (no stray write) and the result must still match the un-guarded run bit for bit (a stray
read of a NaN band would poison it).  Shapes are the ragged ones: row counts and key lengths that are not multiples of the
tile sizes, widths with a partial last vector group."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

GUARD = 32768  # elements (bf16: 64 KB) on each side
SENT = 0x7B7B  # bf16 bit pattern (a finite value nobody computes by accident)


@pytest.fixture(scope="module")
def ops():
    from versecrafter_b200 import _lib, ops
    lib = _lib.load()
    assert lib.vc_device_ok() == 0, lib.vc_last_error()
    return ops


def _rand(shape, seed, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(shape, generator=g, device="cuda") * scale).to(torch.bfloat16)


class Guarded:
    """A [rows, cols] bf16 view with row stride ld >= cols inside a sentinel-filled allocation."""

    def __init__(self, rows, cols, ld=None, src=None, fill_nan=False):
        ld = cols if ld is None else ld
        self.rows, self.cols, self.ld = rows, cols, ld
        n = rows * ld
        self.raw = torch.empty(n + 2 * GUARD, dtype=torch.int16, device="cuda")
        self.fill = 0x7FC0 if fill_nan else SENT  # NaN for inputs, sentinel for outputs
        self.raw.fill_(self.fill)
        self.view = self.raw.view(torch.bfloat16)[GUARD:GUARD + n].view(rows, ld)[:, :cols]
        if src is not None:
            self.view.copy_(src.reshape(rows, cols))

    def intact(self):
        r = self.raw
        body = r[GUARD:GUARD + self.rows * self.ld].view(self.rows, self.ld)
        return bool((r[:GUARD] == self.fill).all()) and bool((r[GUARD + self.rows * self.ld:] == self.fill).all()) and \
            bool((body[:, self.cols:] == self.fill).all())


def gin(t, ld=None):
    """input tensor (2-D) re-housed between NaN guard bands"""
    return Guarded(t.shape[0], t.shape[1], ld, src=t, fill_nan=True)


@pytest.mark.parametrize("M,N,K", [(129, 264, 72), (333, 1536, 200), (77, 64, 128)])
def test_gemm_guard_bands(ops, M, N, K):
    a, w = _rand((M, K), 1), _rand((N, K), 2, 1 / math.sqrt(K))
    bias, gate, res = _rand((N,), 3), _rand((3, N), 4), _rand((M, N), 5)
    rpb = (M + 2) // 3
    cases = [(ops.EPI_BIAS, {}), (ops.EPI_BIAS | ops.EPI_GELU, {}),
             (ops.EPI_BIAS | ops.EPI_GATE | ops.EPI_RES, dict(gate=gate, rows_per_batch=rpb, residual=res)),
             (ops.EPI_BIAS | ops.EPI_SCALE | ops.EPI_RES, dict(residual=res, scale=0.5))]
    for sched in (ops.SCHED_SINGLE, ops.SCHED_PAIR, ops.SCHED_MULTICAST):
        for epi, kw in cases:
            want = ops.gemm(a, w, bias, epilogue=epi | sched, **kw)
            ga, gw, out = gin(a, K + 8), gin(w, K + 16), Guarded(M, N, N + 24)
            kw2 = dict(kw)
            if "residual" in kw2:
                kw2["residual"] = gin(res, N + 8).view
            ops.gemm(ga.view, gw.view, bias, epilogue=epi | sched, out=out.view, **kw2)
            assert out.intact(), (sched, epi)
            assert torch.equal(out.view, want), (sched, epi)


@pytest.mark.parametrize("D", [128, 1032, 1536, 5120, 8192])
def test_row_kernels_guard_bands(ops, D):
    B, rpb = 2, 301
    rows = B * rpb
    x = _rand((rows, D), 6, 2.0)
    m = _rand((B, 6, D), 7, 0.3)
    gamma, beta = _rand((D,), 8), _rand((D,), 9)
    for kw in (dict(), dict(regs=True), dict(ring=True)):
        want = ops.ln_modulate(x, m[:, 0], m[:, 1], rpb, **kw)
        gx, out = gin(x, D + 8), Guarded(rows, D, D + 16)
        ops.ln_modulate(gx.view, m[:, 0], m[:, 1], rpb, out=out.view, **kw)
        assert out.intact() and torch.equal(out.view, want), kw
        want = ops.ln_affine(x, gamma, beta, **kw)
        out = Guarded(rows, D, D + 16)
        ops.ln_affine(gx.view, gamma, beta, out=out.view, **kw)
        assert out.intact() and torch.equal(out.view, want), kw


@pytest.mark.parametrize("D,H", [(256, 2), (1536, 12), (5120, 40)])
def test_rmsnorm_rope_guard_bands(ops, D, H):
    grid, rpb, B = (3, 5, 7), 111, 2   # 105 real positions + 6 pad rows per sample
    rope = ops.rope_table(D // H, grid, "cuda")
    qkv = _rand((B * rpb, 3 * D), 10)
    gq, gk = _rand((D,), 11), _rand((D,), 12)
    for kw in (dict(), dict(regs=True), dict(ring=True)):
        want = ops.rmsnorm_rope_(qkv.clone(), D, 2, gq, gk, rpb, rope=rope, grid=grid, **kw)
        g = Guarded(B * rpb, 3 * D, 3 * D + 8, src=qkv)
        ops.rmsnorm_rope_(g.view, D, 2, gq, gk, rpb, rope=rope, grid=grid, **kw)
        assert g.intact() and torch.equal(g.view, want), kw
        # q only of a cross-attention buffer: the neighbouring columns are not touched
        g1 = Guarded(B * rpb, D, 2 * D, src=qkv[:, :D])
        ops.rmsnorm_rope_(g1.view, D, 1, gq, None, rpb, **kw)
        assert g1.intact(), kw


@pytest.mark.parametrize("variant", [0, 1, 2, 3, 4, 5, 6, 7, 8, 9])
@pytest.mark.parametrize("Lq,Lk", [(333, 333), (129, 515), (37, 5)])
def test_flash_attn_guard_bands(ops, variant, Lq, Lk):
    B, H, d = 2, 3, 128
    q, k, v = _rand((B, Lq, H, d), 13), _rand((B, Lk, H, d), 14), _rand((B, Lk, H, d), 15)
    want = ops.flash_attn(q, k, v, variant=variant)
    gq_, gk_, gv_ = (gin(t.reshape(-1, H * d)) for t in (q, k, v))
    out = Guarded(B * Lq, H * d)
    ops.flash_attn(gq_.view.view(B, Lq, H, d), gk_.view.view(B, Lk, H, d), gv_.view.view(B, Lk, H, d),
                   out=out.view.view(B, Lq, H, d), variant=variant)
    assert out.intact()
    assert torch.equal(out.view.view(B, Lq, H, d), want)


def test_scatter_kernels_guard_bands(ops):
    """Ulysses producers with emulated peers: nothing lands outside the peers' receive blocks."""
    P, H, d, B, K = 4, 8, 128, 2, 192
    grid = (3, 5, 8)
    L = 120
    Lp, Hp, D = L // P, H // P, H * d
    rope = ops.rope_table(d, grid, "cuda")
    gq, gk = _rand((D,), 16), _rand((D,), 17)
    h = _rand((B * Lp, K), 18)
    w, bias = _rand((3 * D, K), 19, 1 / math.sqrt(K)), _rand((3 * D,), 20)
    qkv = ops.gemm(h, w, bias)
    recv = [Guarded(3 * B * L * Hp, d) for _ in range(P)]
    attn = [Guarded(B * Lp, D) for _ in range(P)]
    r = 2
    ptrs = [g.view.data_ptr() for g in recv]
    ops.rmsnorm_rope_qkv_scatter(qkv, D, gq, gk, Lp, ptrs, Hp, L, rope, grid, pos_offset=r * Lp, which_mask=1)
    ops.rmsnorm_rope_qkv_scatter(qkv, D, gq, gk, Lp, ptrs, Hp, L, rope, grid, pos_offset=r * Lp, which_mask=2, bulk=True)
    ops.gemm_scatter(h, w[2 * D:], bias[2 * D:], Lp, [p + 2 * B * L * Hp * d * 2 for p in ptrs], Hp, L, pos_offset=r * Lp)
    assert all(g.intact() for g in recv)
    # rows of rank r are written in all three matrices, every other row still holds the sentinel
    body = recv[0].raw[GUARD:GUARD + 3 * B * L * Hp * d].view(3, B, L, Hp * d)
    assert bool((body[:, :, r * Lp:(r + 1) * Lp] != SENT).any(dim=-1).all())
    assert bool((body[:, :, :r * Lp] == SENT).all()) and bool((body[:, :, (r + 1) * Lp:] == SENT).all())
    qf, kf, vf = (_rand((B, L, Hp, d), 21 + i) for i in range(3))
    ops.flash_attn_scatter(qf, kf, vf, [g.view.data_ptr() + 1 * Hp * d * 2 for g in attn], Lp, (Lp * D, D, d), variant=4)
    assert all(g.intact() for g in attn)
    a0 = attn[0].raw[GUARD:GUARD + B * Lp * D].view(B * Lp, H, d)
    assert bool((a0[:, Hp:2 * Hp] != SENT).any(dim=-1).all())                       # this "rank"'s heads arrived
    assert bool((a0[:, :Hp] == SENT).all()) and bool((a0[:, 2 * Hp:] == SENT).all())  # the other heads' slots untouched


@pytest.mark.parametrize("C,Fr,Hh,Ww,D", [(128, 3, 10, 24, 264), (16, 2, 12, 104, 256)])
def test_patch_embed_guard_bands(ops, C, Fr, Hh, Ww, D):
    u = _rand((C, Fr, Hh, Ww), 30)
    w2d, bias = _rand((D, 4 * C), 31, 0.05), _rand((D,), 32)
    L = Fr * (Hh // 2) * (Ww // 2)
    lo, hi = 7, L - 5
    want = torch.empty((hi - lo, D), dtype=torch.bfloat16, device="cuda")
    ops.patch_embed(u, w2d, bias, want, lo, hi)
    gu = Guarded(1, u.numel(), src=u.reshape(1, -1), fill_nan=True)
    out = Guarded(hi - lo, D, D + 8)
    ops.patch_embed(gu.view.view(C, Fr, Hh, Ww), w2d, bias, out.view, lo, hi)
    assert out.intact() and torch.equal(out.view, want)
    y = _rand((L, 64), 33)
    pat = ops.patchify(u)
    assert pat.shape == (L, 4 * C)
    un = ops.unpatchify(y, (Fr, Hh // 2, Ww // 2))
    assert un.shape == (16, Fr, Hh, Ww)
