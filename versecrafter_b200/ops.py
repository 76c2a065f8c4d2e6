"""Thin torch-tensor wrappers over the C ABI (include/vc_b200.h).  torch is used for device memory and the current
stream only; every op below is one launch of a hand-written sm_100a kernel."""
import ctypes
import math

import torch

from . import _lib

EPI_BIAS, EPI_GELU, EPI_GATE, EPI_RES, EPI_SCALE = 1, 2, 4, 8, 16
SCHED_MULTICAST = 64   # schedule bits of vc_gemm_bf16 (same arithmetic): cluster of 2 sharing W by TMA multicast,
SCHED_PAIR = 128       # CTA pair issuing one cta_group::2 MMA of M = 256,
SCHED_SINGLE = 256     # one CTA per tile (overrides the VC_GEMM_SCHED default)

# kernels launched through this module since the last reset (bench.py reports it as gpu_launches)
launch_count = 0

# optional per-kernel-class device timing: bench.py sets `timing = {}`; every wrapper then brackets its launch with
# CUDA events on the current stream and appends (start, end, work) under its class name.
timing = None


class _timed:
    def __init__(self, name, work=0.0):
        self.name, self.work = name, work

    def __enter__(self):
        if timing is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *a):
        if timing is not None:
            self.e1.record()
            timing.setdefault(self.name, []).append((self.e0, self.e1, self.work))
        return False


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def _chk(t, name, dtype=torch.bfloat16):
    if not t.is_cuda:
        raise _lib.VCError(f"{name}: expected a CUDA tensor (the sm_100a path has no CPU fallback)")
    if t.dtype != dtype:
        raise _lib.VCError(f"{name}: expected {dtype}, got {t.dtype}")
    if t.stride(-1) != 1:
        raise _lib.VCError(f"{name}: innermost dimension must be contiguous")


def _rows2d(t, name):
    """View [..., D] as rows with one row stride; returns (rows, ld)."""
    _chk(t, name)
    if t.dim() == 2:
        return t.shape[0], t.stride(0)
    tt = t if t.is_contiguous() else None
    if tt is None:
        raise _lib.VCError(f"{name}: >2-D tensors must be contiguous")
    return t.numel() // t.shape[-1], t.shape[-1]


def ln_modulate(x, shift, scale, rows_per_batch, eps=1e-6, out=None, regs=False, ring=False):
    """out = LN(x) * (1 + scale[b]) + shift[b];  x [rows, D] (or [B, L, D]); shift/scale [B, D] views (bf16)."""
    global launch_count
    rows, ldx = _rows2d(x, "x")
    D = x.shape[-1]
    _chk(shift, "shift"); _chk(scale, "scale")
    assert shift.shape == scale.shape and shift.shape[-1] == D
    sh2, sc2 = shift.reshape(-1, D) if shift.is_contiguous() else shift, scale.reshape(-1, D) if scale.is_contiguous() else scale
    assert sh2.dim() == 2 and sc2.dim() == 2 and sh2.stride(0) == sc2.stride(0)
    if out is None:
        out = torch.empty(x.shape, dtype=x.dtype, device=x.device)
    _, ldo = _rows2d(out, "out")
    lib = _lib.load()
    with _timed("ln_modulate", 4.0 * rows * D):
        _lib.check(lib.vc_ln_modulate(_ptr(x), ldx, _ptr(out), ldo, _ptr(sh2), _ptr(sc2), sh2.stride(0), rows,
                                      rows_per_batch, D, eps, (16 if regs else 0) | (32 if ring else 0), _stream()), "vc_ln_modulate")
    launch_count += 1
    return out


def ln_affine(x, gamma, beta, eps=1e-6, out=None, regs=False, ring=False):
    """out = LN(x) * gamma + beta (norm3)."""
    global launch_count
    rows, ldx = _rows2d(x, "x")
    D = x.shape[-1]
    _chk(gamma, "gamma"); _chk(beta, "beta")
    if out is None:
        out = torch.empty(x.shape, dtype=x.dtype, device=x.device)
    _, ldo = _rows2d(out, "out")
    lib = _lib.load()
    with _timed("ln_affine", 4.0 * rows * D):
        _lib.check(lib.vc_ln_modulate(_ptr(x), ldx, _ptr(out), ldo, _ptr(beta), _ptr(gamma), 0, rows, max(rows, 1), D, eps,
                                      1 | (16 if regs else 0) | (32 if ring else 0), _stream()), "vc_ln_modulate")
    launch_count += 1
    return out


def rope_table(head_dim, grid, device, theta=10000.0):
    """(cos, sin) float2 tables per axis, fp64 angles rounded to fp32 (rope_params, wan_transformer3d.py:52-60 and the
    table split at :783-795).  Returns (table [F*nf + H*nh + W*nw, 2] fp32, (nf, nh, nw))."""
    c = head_dim // 2
    nh = nw = c // 3
    nf = c - 2 * (c // 3)
    parts = []
    for n_pos, n in zip(grid, (nf, nh, nw)):
        dim = 2 * n
        inv = 1.0 / torch.pow(torch.tensor(theta, dtype=torch.float64),
                              torch.arange(0, dim, 2, dtype=torch.float64) / dim)
        ang = torch.arange(n_pos, dtype=torch.float64)[:, None] * inv[None]
        parts.append(torch.stack([ang.cos(), ang.sin()], dim=-1).reshape(-1, 2))
    return torch.cat(parts).to(torch.float32).to(device).contiguous(), (nf, nh, nw)


def rmsnorm_rope_(buf, D, n_mats, gain_q, gain_k, rows_per_batch, eps=1e-6, rope=None, grid=None, pos_offset=0, regs=False, ring=False):
    """In place: first n_mats D-wide column blocks of buf [rows, ld] <- RoPE(RMSNorm(.) * gain).
    rope = (table, (nf, nh, nw)) from rope_table, grid = (F, H', W'); rope=None -> RMSNorm only."""
    global launch_count
    _chk(buf, "buf")
    assert buf.dim() == 2
    rows, ld = buf.shape[0], buf.stride(0)
    _chk(gain_q, "gain_q")
    if gain_k is not None:
        _chk(gain_k, "gain_k")
    lib = _lib.load()
    if rope is not None:
        table, (nf, nh, nw) = rope
        gf, gh, gw = grid
        tptr, do = _ptr(table), 1
    else:
        nf = nh = nw = 0
        gf = gh = gw = 1
        tptr, do = ctypes.c_void_p(0), 0
    with _timed("rmsnorm_rope", 4.0 * rows * D * n_mats):
        _lib.check(lib.vc_rmsnorm_rope_qk(_ptr(buf), ld, rows, rows_per_batch, pos_offset, D, n_mats, _ptr(gain_q),
                                          _ptr(gain_k), tptr, gf, gh, gw, nf, nh, nw, eps, do | (16 if regs else 0) | (32 if ring else 0), _stream()),
                   "vc_rmsnorm_rope_qk")
    launch_count += 1
    return buf


def time_embed(t, w0, b0, w2, b2, wp, bp, e_out=None, e0_out=None, workspace=None):
    """fp32 timestep-embedding island: t [B] (any real dtype) -> (e [B, D] bf16, e0 [B, 6, D] bf16).  Four launches
    (fp64 sinusoid + three weight-streaming GEMVs); w0 [D, freq_dim], w2 [D, D], wp [6D, D] bf16."""
    global launch_count
    for u, n in ((w0, "w0"), (b0, "b0"), (w2, "w2"), (b2, "b2"), (wp, "wp"), (bp, "bp")):
        _chk(u, n)
        if not u.is_contiguous():
            raise _lib.VCError(f"time_embed: {n} must be contiguous")
    if not t.is_cuda:
        raise _lib.VCError("time_embed: t must be a CUDA tensor")
    B, (D, freq_dim) = t.numel(), w0.shape
    t64 = t.reshape(-1).to(torch.float64)
    dev = w0.device
    if e_out is None:
        e_out = torch.empty((B, D), dtype=torch.bfloat16, device=dev)
    if e0_out is None:
        e0_out = torch.empty((B, 6, D), dtype=torch.bfloat16, device=dev)
    if workspace is None:
        workspace = torch.empty(B * (freq_dim + 2 * D), dtype=torch.float32, device=dev)
    assert workspace.dtype == torch.float32 and workspace.numel() >= B * (freq_dim + 2 * D)
    lib = _lib.load()
    with _timed("time_embed", 2.0 * (D * freq_dim + 7 * D * D)):
        _lib.check(lib.vc_time_embed(_ptr(t64), B, freq_dim, D, _ptr(w0), _ptr(b0), _ptr(w2), _ptr(b2), _ptr(wp), _ptr(bp),
                                     _ptr(e_out), _ptr(e0_out), _ptr(workspace), _stream()), "vc_time_embed")
    launch_count += 4
    return e_out, e0_out


def modulation_table(mod_ptrs, n_blocks, e0, out, head_mod=None, e=None, out_head=None):
    """out[i, b] = bf16(modulation_i + e0[b]) for every block in one launch; mod_ptrs: int64 CUDA tensor holding the
    device addresses of the blocks' [1, 6, D] modulation parameters.  Optionally the head's [B, 2, D] table from e."""
    global launch_count
    _chk(e0, "e0"); _chk(out, "out")
    B, D = e0.shape[0], e0.shape[-1]
    assert e0.is_contiguous() and out.is_contiguous() and out.numel() >= n_blocks * B * 6 * D
    assert mod_ptrs.dtype == torch.int64 and mod_ptrs.is_cuda and mod_ptrs.numel() >= n_blocks
    if head_mod is not None:
        _chk(head_mod, "head_mod"); _chk(e, "e"); _chk(out_head, "out_head")
        assert e.is_contiguous() and out_head.is_contiguous() and head_mod.is_contiguous()
    lib = _lib.load()
    _lib.check(lib.vc_modulation_table(_ptr(mod_ptrs), n_blocks, _ptr(e0), B, D, _ptr(out), _ptr(head_mod), _ptr(e),
                                       _ptr(out_head), _stream()), "vc_modulation_table")
    launch_count += 1
    return out


def gemm(a, w, bias=None, epilogue=None, gate=None, rows_per_batch=0, residual=None, scale=1.0, out=None):
    """out[M,N] = epilogue(a[M,K] @ w[N,K]^T).  a / out / residual may be row-strided 2-D views."""
    global launch_count
    _chk(a, "a"); _chk(w, "w")
    assert a.dim() == 2 and w.dim() == 2 and a.shape[1] == w.shape[1], (a.shape, w.shape)
    M, K = a.shape
    N = w.shape[0]
    if epilogue is None:
        epilogue = EPI_BIAS if bias is not None else 0
    if out is None:
        out = torch.empty((M, N), dtype=torch.bfloat16, device=a.device)
    _chk(out, "out")
    assert out.shape == (M, N)
    gstride = 0
    if gate is not None:
        _chk(gate, "gate")
        assert gate.dim() == 2 and gate.shape[1] == N
        gstride = gate.stride(0)
    ldr = 0
    if residual is not None:
        _chk(residual, "residual")
        assert residual.shape == (M, N)
        ldr = residual.stride(0)
    if bias is not None:
        _chk(bias, "bias")
    lib = _lib.load()
    with _timed("gemm", 2.0 * M * N * K):
        _lib.check(lib.vc_gemm_bf16(_ptr(a), a.stride(0), _ptr(w), w.stride(0), _ptr(out), out.stride(0), M, N, K,
                                    epilogue, _ptr(bias), _ptr(gate), gstride, rows_per_batch, _ptr(residual), ldr,
                                    float(scale), _stream()), "vc_gemm_bf16")
    launch_count += 1
    return out


def _strides3(t):
    # t: [B, L, H, d] view
    return (ctypes.c_longlong * 3)(t.stride(0), t.stride(1), t.stride(2))


def flash_attn(q, k, v, out=None, softmax_scale=None, variant=4):
    """q [B, Lq, H, 128], k/v [B, Lk, H, 128] (arbitrary batch/token/head strides, unit stride on head_dim)
    -> out [B, Lq, H, 128]."""
    global launch_count
    for t, n in ((q, "q"), (k, "k"), (v, "v")):
        _chk(t, n)
        assert t.dim() == 4 and t.shape[-1] == 128, (n, t.shape)
    B, Lq, H, d = q.shape
    Lk = k.shape[1]
    assert k.shape == (B, Lk, H, d) and v.shape == (B, Lk, H, d)
    if out is None:
        out = torch.empty((B, Lq, H, d), dtype=torch.bfloat16, device=q.device)
    _chk(out, "out")
    if softmax_scale is None:
        softmax_scale = 1.0 / math.sqrt(d)
    lib = _lib.load()
    with _timed("attn_self" if Lk > 1024 else "attn_cross", 4.0 * B * H * Lq * Lk * d):
        _lib.check(lib.vc_flash_attn_fwd(_ptr(q), _ptr(k), _ptr(v), _ptr(out), B, H, Lq, Lk, d, _strides3(q),
                                         _strides3(k), _strides3(v), _strides3(out), float(softmax_scale), int(variant),
                                         _stream()), "vc_flash_attn_fwd")
    launch_count += 1
    return out


def _ptr_array(ptrs):
    return (ctypes.c_void_p * len(ptrs))(*[int(x) for x in ptrs])


def rmsnorm_rope_qkv_scatter(qkv, D, gain_q, gain_k, rows_per_batch, peer_ptrs, heads_per_peer, l_total, rope, grid,
                             pos_offset=0, eps=1e-6, which_mask=7, bulk=False, cta_cap=0):
    """qkv [rows, 3D] of this rank's tokens -> RMSNorm+RoPE(q,k), v, written head-scattered into the peers' receive
    buffers (device addresses `peer_ptrs`, layout [3, B, l_total, heads_per_peer, d]).  which_mask: bit 0 q, 1 k, 2 v.
    bulk: stores issued by the TMA engine (one cp.async.bulk per row, matrix and peer); cta_cap: resident CTAs per SM."""
    global launch_count
    _chk(qkv, "qkv"); _chk(gain_q, "gain_q"); _chk(gain_k, "gain_k")
    assert qkv.dim() == 2 and qkv.shape[1] >= 3 * D
    rows, ld = qkv.shape[0], qkv.stride(0)
    table, (nf, nh, nw) = rope
    gf, gh, gw = grid
    lib = _lib.load()
    with _timed("rmsnorm_rope", 4.0 * bin(which_mask & 7).count("1") * rows * D):
        _lib.check(lib.vc_rmsnorm_rope_qkv_scatter(_ptr(qkv), ld, rows, rows_per_batch, pos_offset, D, _ptr(gain_q),
                                                   _ptr(gain_k), _ptr(table), gf, gh, gw, nf, nh, nw, eps,
                                                   _ptr_array(peer_ptrs), len(peer_ptrs), heads_per_peer, l_total,
                                                   int(which_mask) | (8 if bulk else 0) | (int(cta_cap) << 8), _stream()),
                   "vc_rmsnorm_rope_qkv_scatter")
    launch_count += 1


def gemm_scatter(a, w, bias, rows_per_batch, peer_ptrs, heads_per_peer, l_total, pos_offset=0, head_dim=128, sched=0):
    """(a[M,K] @ w[N,K]^T + bias) with head h of row r stored into peer h // heads_per_peer's receive block
    [n_batch, l_total, heads_per_peer, head_dim] (device addresses `peer_ptrs`) at token pos_offset + r % rows_per_batch:
    a Linear and its Ulysses head-scatter in one kernel."""
    global launch_count
    _chk(a, "a"); _chk(w, "w"); _chk(bias, "bias")
    assert a.dim() == 2 and w.dim() == 2 and a.shape[1] == w.shape[1], (a.shape, w.shape)
    M, K = a.shape
    N = w.shape[0]
    lib = _lib.load()
    with _timed("gemm", 2.0 * M * N * K):
        _lib.check(lib.vc_gemm_bf16_scatter(_ptr(a), a.stride(0), _ptr(w), w.stride(0), M, N, K, _ptr(bias), rows_per_batch,
                                            pos_offset, _ptr_array(peer_ptrs), len(peer_ptrs), heads_per_peer, head_dim,
                                            l_total, int(sched), _stream()), "vc_gemm_bf16_scatter")
    launch_count += 1


def flash_attn_scatter(q, k, v, peer_ptrs, rows_per_peer, o_strides, softmax_scale=None, variant=4):
    """q,k,v [B, L, Hp, 128] (local heads, all tokens); output row t -> row t % rows_per_peer of peer t // rows_per_peer
    (device addresses `peer_ptrs`, element strides o_strides = (batch, token, head))."""
    global launch_count
    for t, n in ((q, "q"), (k, "k"), (v, "v")):
        _chk(t, n)
        assert t.dim() == 4 and t.shape[-1] == 128, (n, t.shape)
    B, Lq, H, d = q.shape
    Lk = k.shape[1]
    if softmax_scale is None:
        softmax_scale = 1.0 / math.sqrt(d)
    lib = _lib.load()
    with _timed("attn_self" if Lk > 1024 else "attn_cross", 4.0 * B * H * Lq * Lk * d):
        _lib.check(lib.vc_flash_attn_fwd_scatter(_ptr(q), _ptr(k), _ptr(v), B, H, Lq, Lk, d, _strides3(q), _strides3(k),
                                                 _strides3(v), _ptr_array(peer_ptrs), len(peer_ptrs), rows_per_peer,
                                                 (ctypes.c_longlong * 3)(*o_strides), float(softmax_scale), int(variant),
                                                 _stream()), "vc_flash_attn_fwd_scatter")
    launch_count += 1


def patch_embed_supported(C, Fr, H, W, D):
    return bool(_lib.load().vc_patch_embed_supported(C, Fr, H, W, D))


def patch_embed(u, weight2d, bias, out, tok_lo=0, tok_hi=None):
    """Conv3d(kernel = stride = (1,2,2)) of one sample u [C, F, H, W] as a TMA-fed implicit GEMM:
    out[n - tok_lo] = conv(u)[:, n] + bias for tokens n in [tok_lo, tok_hi); weight2d = conv.weight.view(D, 4C)."""
    global launch_count
    _chk(u, "u"); _chk(weight2d, "weight"); _chk(bias, "bias"); _chk(out, "out")
    assert u.dim() == 4 and u.is_contiguous() and weight2d.dim() == 2 and out.dim() == 2
    C, Fr, H, W = u.shape
    D = weight2d.shape[0]
    L = Fr * (H // 2) * (W // 2)
    tok_hi = L if tok_hi is None else tok_hi
    assert weight2d.shape[1] == 4 * C and out.shape[1] == D and out.shape[0] >= tok_hi - tok_lo
    lib = _lib.load()
    with _timed("patch_embed", 2.0 * (tok_hi - tok_lo) * D * 4 * C):
        _lib.check(lib.vc_patch_embed(_ptr(u), C, Fr, H, W, _ptr(weight2d), weight2d.stride(0), _ptr(bias), D, _ptr(out),
                                      out.stride(0), tok_lo, tok_hi, _stream()), "vc_patch_embed")
    launch_count += 1
    return out


def patchify(u, out=None):
    """u [C, F, H, W] bf16 -> [L, 4C]."""
    global launch_count
    _chk(u, "u")
    assert u.dim() == 4 and u.is_contiguous()
    C, Fr, H, W = u.shape
    L = Fr * (H // 2) * (W // 2)
    if out is None:
        out = torch.empty((L, 4 * C), dtype=torch.bfloat16, device=u.device)
    lib = _lib.load()
    _lib.check(lib.vc_patchify(_ptr(u), _ptr(out), C, Fr, H, W, out.stride(0), _stream()), "vc_patchify")
    launch_count += 1
    return out


def unpatchify(y, grid, out_dim=16):
    """y [L, 4*out_dim] -> [out_dim, F, 2H', 2W']."""
    global launch_count
    _chk(y, "y")
    gf, gh, gw = grid
    out = torch.empty((out_dim, gf, 2 * gh, 2 * gw), dtype=torch.bfloat16, device=y.device)
    lib = _lib.load()
    _lib.check(lib.vc_unpatchify(_ptr(y), y.stride(0), _ptr(out), out_dim, gf, gh, gw, _stream()), "vc_unpatchify")
    launch_count += 1
    return out
