"""Per-kernel timing on one B200 (CUDA events, warm-up, L2 flushed between iterations by working sets > L2).
Writes JSON lines to gpurun_out/microbench.jsonl.  Library baselines (torch.matmul = cuBLASLt, flash_attn 2.8) are
timed beside each kernel as the bar to beat; they are not on the product path."""
import json
import math
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from versecrafter_b200 import ops  # noqa: E402

PEAKS = {"hbm_gbs": 6548.5, "bf16_tflops": 1604.8, "bf16_tflops_sustained": 1381.2}
try:
    PEAKS.update(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))))
except Exception:
    pass
OUT = os.path.join(os.path.dirname(__file__), "..", "gpurun_out", "microbench.jsonl")


def timeit(fn, iters=10, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = [ev[i].elapsed_time(ev[i + 1]) for i in range(iters)]
    return sum(ts) / len(ts), min(ts)


def emit(rec):
    print(json.dumps(rec), flush=True)
    with open(OUT, "a") as f:
        f.write(json.dumps(rec) + "\n")


def rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


def bench_rows(B=2, L=32760, D=5120):
    x = torch.randn(B, L, D, device="cuda", dtype=torch.bfloat16)
    m = torch.randn(B, 6, D, device="cuda", dtype=torch.bfloat16) * 0.1
    out = torch.empty_like(x)
    avg, mn = timeit(lambda: ops.ln_modulate(x, m[:, 0], m[:, 1], L, out=out))
    bytes_ = 2 * x.numel() * 2
    emit(dict(kernel="ln_modulate", shape=[B, L, D], ms=avg, ms_min=mn, gbs=bytes_ / avg / 1e6,
              frac_hbm=bytes_ / avg / 1e6 / PEAKS["hbm_gbs"]))
    gam = torch.randn(D, device="cuda", dtype=torch.bfloat16)
    avg, mn = timeit(lambda: ops.ln_affine(x, gam, gam, out=out))
    emit(dict(kernel="ln_affine", shape=[B, L, D], ms=avg, ms_min=mn, gbs=bytes_ / avg / 1e6,
              frac_hbm=bytes_ / avg / 1e6 / PEAKS["hbm_gbs"]))
    del out
    qkv = torch.randn(B * L, 3 * D, device="cuda", dtype=torch.bfloat16)
    g = torch.ones(D, device="cuda", dtype=torch.bfloat16)
    grid = (21, 30, 52) if L == 32760 else (1, 1, L)
    rope = ops.rope_table(128, grid, "cuda")
    avg, mn = timeit(lambda: ops.rmsnorm_rope_(qkv, D, 2, g, g, L, rope=rope, grid=grid))
    bytes_ = 2 * 2 * B * L * D * 2
    emit(dict(kernel="rmsnorm_rope_qk", shape=[B * L, 3 * D], ms=avg, ms_min=mn, gbs=bytes_ / avg / 1e6,
              frac_hbm=bytes_ / avg / 1e6 / PEAKS["hbm_gbs"]))


def bench_rows_ab(B=2, L=32760, D=5120, rounds=4):
    """Row kernels at the c3 launch shape: asynchronous row ring (round 2) vs register-resident (round 1), interleaved."""
    x = torch.randn(B, L, D, device="cuda", dtype=torch.bfloat16)
    m = torch.randn(B, 6, D, device="cuda", dtype=torch.bfloat16) * 0.1
    out = torch.empty_like(x)
    gam = torch.randn(D, device="cuda", dtype=torch.bfloat16)
    qkv = torch.randn(B * L, 3 * D, device="cuda", dtype=torch.bfloat16)
    g = torch.ones(D, device="cuda", dtype=torch.bfloat16)
    grid = (21, 30, 52) if L == 32760 else (1, 1, L)
    rope = ops.rope_table(128, grid, "cuda")
    qc = torch.randn(B * L, D, device="cuda", dtype=torch.bfloat16)
    cases = {
        "ln_modulate": (lambda r: ops.ln_modulate(x, m[:, 0], m[:, 1], L, out=out, regs=r, ring=not r), 4.0 * B * L * D),
        "ln_affine": (lambda r: ops.ln_affine(x, gam, gam, out=out, regs=r, ring=not r), 4.0 * B * L * D),
        "rmsnorm_rope_qk": (lambda r: ops.rmsnorm_rope_(qkv, D, 2, g, g, L, rope=rope, grid=grid, regs=r, ring=not r), 8.0 * B * L * D),
        "rmsnorm_q_cross": (lambda r: ops.rmsnorm_rope_(qc, D, 1, g, None, L, regs=r, ring=not r), 4.0 * B * L * D),
    }
    for name, (fn, bytes_) in cases.items():
        t = {"ring": [], "regs": []}
        for _ in range(rounds):
            for key, r in (("ring", False), ("regs", True)):
                avg, _ = timeit(lambda: fn(r), iters=10, warmup=2)
                t[key].append(avg)
        rec = dict(kernel=name + "_ab", shape=[B, L, D])
        for key, ts in t.items():
            ms = sum(ts) / len(ts)
            rec[key] = dict(ms=ms, gbs=bytes_ / ms / 1e6, frac_hbm=bytes_ / ms / 1e6 / PEAKS["hbm_gbs"])
        emit(rec)


def bench_gemm(M, N, K, epi="bias"):
    a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16) / math.sqrt(K)
    b = torch.randn(N, device="cuda", dtype=torch.bfloat16)
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    kw = {}
    if epi == "gelu":
        kw = dict(epilogue=3)
    elif epi == "gate_res":
        gate = torch.randn(2, N, device="cuda", dtype=torch.bfloat16)
        res = torch.randn(M, N, device="cuda", dtype=torch.bfloat16)
        kw = dict(epilogue=13, gate=gate, rows_per_batch=(M + 1) // 2, residual=res)
    avg, mn = timeit(lambda: ops.gemm(a, w, b, out=out, **kw), iters=6)
    flops = 2.0 * M * N * K
    ref_avg, ref_mn = timeit(lambda: torch.nn.functional.linear(a, w, b), iters=6)
    err = rel(ops.gemm(a[:4096], w, b), torch.nn.functional.linear(a[:4096], w, b)) if epi == "bias" else None
    emit(dict(kernel="gemm_" + epi, shape=[M, N, K], ms=avg, ms_min=mn, tflops=flops / avg / 1e9,
              frac_bf16_sustained=flops / avg / 1e9 / PEAKS["bf16_tflops_sustained"],
              frac_bf16_burst=flops / mn / 1e9 / PEAKS["bf16_tflops"],
              cublas_ms=ref_avg, cublas_tflops=flops / ref_avg / 1e9, rel_err_vs_cublas=err))


def bench_gemm_sched(M, N, K, epi="bias", rounds=4, iters=4):
    """The three schedules of vc_gemm_bf16 (one CTA per tile / multicast cluster / cta_group::2 pair) and cuBLASLt on one
    shape, interleaved round-robin so they see the same thermal / power state."""
    a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16) / math.sqrt(K)
    b = torch.randn(N, device="cuda", dtype=torch.bfloat16)
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    base, kw = 1, {}
    if epi == "gelu":
        base = 3
    elif epi == "gate_res":
        gate = torch.randn(2, N, device="cuda", dtype=torch.bfloat16)
        res = torch.randn(M, N, device="cuda", dtype=torch.bfloat16)
        base, kw = 13, dict(gate=gate, rows_per_batch=(M + 1) // 2, residual=res)
    arms = {"single": lambda: ops.gemm(a, w, b, out=out, epilogue=base | ops.SCHED_SINGLE, **kw),
            "mc": lambda: ops.gemm(a, w, b, out=out, epilogue=base | ops.SCHED_MULTICAST, **kw),
            "pair": lambda: ops.gemm(a, w, b, out=out, epilogue=base | ops.SCHED_PAIR, **kw),
            "cublaslt": lambda: torch.nn.functional.linear(a, w, b)}
    ref = arms["single"]().clone()
    same = {k: bool(torch.equal(arms[k](), ref)) for k in ("mc", "pair")}
    times = {k: [] for k in arms}
    for _ in range(rounds):
        for k, fn in arms.items():
            avg, _ = timeit(fn, iters=iters, warmup=1)
            times[k].append(avg)
    flops = 2.0 * M * N * K
    rec = dict(kernel="gemm_sched_" + epi, shape=[M, N, K], bit_identical=same)
    for k, ts in times.items():
        ms = sum(ts) / len(ts)
        rec[k] = dict(ms=ms, ms_min=min(ts), tflops=flops / ms / 1e9)
    emit(rec)


def bench_attn(B, H, Lq, Lk, variant):
    q = torch.randn(B, Lq, H, 128, device="cuda", dtype=torch.bfloat16)
    k = torch.randn(B, Lk, H, 128, device="cuda", dtype=torch.bfloat16)
    v = torch.randn(B, Lk, H, 128, device="cuda", dtype=torch.bfloat16)
    out = torch.empty_like(q)
    avg, mn = timeit(lambda: ops.flash_attn(q, k, v, out=out, variant=variant), iters=5, warmup=2)
    flops = 4.0 * B * H * Lq * Lk * 128
    rec = dict(kernel=f"flash_attn_v{variant}", shape=[B, H, Lq, Lk], ms=avg, ms_min=mn, tflops=flops / avg / 1e9,
               frac_bf16_sustained=flops / avg / 1e9 / PEAKS["bf16_tflops_sustained"])
    try:
        from flash_attn import flash_attn_func
        ref = flash_attn_func(q, k, v)
        rec["rel_err_vs_fa2"] = rel(out, ref)
        ravg, _ = timeit(lambda: flash_attn_func(q, k, v), iters=5, warmup=2)
        rec["fa2_ms"] = ravg
        rec["fa2_tflops"] = flops / ravg / 1e9
    except Exception as e:  # noqa: BLE001
        rec["fa2_error"] = repr(e)[:200]
    emit(rec)


def bench_attn_sustained(variant, seconds=4.0, B=2, H=40, L=32760):
    """c3 self-attention launched back to back for a few seconds: throughput under the 1 kW power cap (what the
    kernel delivers inside a long denoise step), not the burst figure."""
    q = torch.randn(B, L, H, 128, device="cuda", dtype=torch.bfloat16)
    k = torch.randn(B, L, H, 128, device="cuda", dtype=torch.bfloat16)
    v = torch.randn(B, L, H, 128, device="cuda", dtype=torch.bfloat16)
    out = torch.empty_like(q)
    ops.flash_attn(q, k, v, out=out, variant=variant)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ops.flash_attn(q, k, v, out=out, variant=variant)
    e1.record()
    torch.cuda.synchronize()
    n = max(4, int(seconds * 1e3 / e0.elapsed_time(e1)))
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
    ev[0].record()
    for i in range(n):
        ops.flash_attn(q, k, v, out=out, variant=variant)
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = [ev[i].elapsed_time(ev[i + 1]) for i in range(n)]
    flops = 4.0 * B * H * L * L * 128
    last = ts[len(ts) // 2:]
    emit(dict(kernel=f"flash_attn_v{variant}_sustained", shape=[B, H, L, L], launches=n, ms_first=ts[0],
              ms_last_half=sum(last) / len(last), tflops_first=flops / ts[0] / 1e9,
              tflops_last_half=flops / (sum(last) / len(last)) / 1e9,
              frac_bf16_sustained=flops / (sum(last) / len(last)) / 1e9 / PEAKS["bf16_tflops_sustained"]))


def bench_patch_embed(C=128, grid=(21, 60, 104), D=5120):
    """Control-map patch embedding (K2) at the c3 shape: implicit GEMM vs patchify + GEMM vs torch conv3d (cuDNN)."""
    Fr, H, W = grid
    u = torch.randn(C, Fr, H, W, device="cuda", dtype=torch.bfloat16)
    wt = torch.randn(D, C, 1, 2, 2, device="cuda", dtype=torch.bfloat16) / math.sqrt(4 * C)
    b = torch.randn(D, device="cuda", dtype=torch.bfloat16)
    L = Fr * (H // 2) * (W // 2)
    out = torch.empty(L, D, device="cuda", dtype=torch.bfloat16)
    A = torch.empty(L, 4 * C, device="cuda", dtype=torch.bfloat16)
    avg, mn = timeit(lambda: ops.patch_embed(u, wt.view(D, -1), b, out))
    ex_avg, _ = timeit(lambda: ops.gemm(ops.patchify(u, out=A), wt.view(D, -1), b, out=out))
    cd_avg, _ = timeit(lambda: torch.nn.functional.conv3d(u[None], wt, b, stride=(1, 2, 2)).flatten(2).transpose(1, 2).contiguous())
    flops = 2.0 * L * D * 4 * C
    bytes_ = (u.numel() + out.numel() + wt.numel()) * 2
    emit(dict(kernel="patch_embed_implicit", shape=[C, Fr, H, W, D], ms=avg, ms_min=mn, tflops=flops / avg / 1e9,
              gbs=bytes_ / avg / 1e6, frac_hbm=bytes_ / avg / 1e6 / PEAKS["hbm_gbs"],
              frac_bf16_sustained=flops / avg / 1e9 / PEAKS["bf16_tflops_sustained"], explicit_ms=ex_avg, cudnn_conv3d_ms=cd_avg))


if __name__ == "__main__":
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    which = sys.argv[1:] or ["rows", "gemm", "attn"]
    torch.cuda.init()
    if "rows" in which:
        bench_rows()
    if "gemm" in which:
        M = 65520
        for (N, K, epi) in [(15360, 5120, "bias"), (5120, 5120, "gate_res"), (13824, 5120, "gelu"), (5120, 13824, "gate_res"),
                            (5120, 512, "bias"), (4608, 1536, "bias")]:
            bench_gemm(M, N, K, epi)
    if "rows_ab" in which:
        bench_rows_ab()
        bench_rows_ab(D=1536)
    if "gemm_sched" in which and os.environ.get("VC_MB_GEMM_ONE"):   # one shape only (ncu capture of the pair kernel)
        bench_gemm_sched(65520, 15360, 5120, "bias")
    elif "gemm_sched" in which:
        for (M, N, K, epi) in [(65520, 15360, 5120, "bias"), (65520, 5120, 5120, "gate_res"), (65520, 13824, 5120, "gelu"),
                               (65520, 5120, 13824, "gate_res"), (65520, 4608, 1536, "bias"), (65520, 8960, 1536, "gelu"),
                               (65520, 1536, 8960, "gate_res"), (8190, 15360, 5120, "bias"), (8190, 5120, 13824, "gate_res")]:
            bench_gemm_sched(M, N, K, epi)
    if "attn_sustained" in which:
        for variant in [int(v) for v in os.environ.get("VC_ATTN_VARIANTS", "1,2").split(",")]:
            bench_attn_sustained(variant)
    if "patch_embed" in which:
        bench_patch_embed()
        bench_patch_embed(C=16)
    if "attn_c3" in which:  # the per-launch shape of the c3 step (B=2, 40 heads, 32,760 tokens)
        for variant in [int(v) for v in os.environ.get("VC_ATTN_VARIANTS", "4").split(",")]:
            bench_attn(2, 40, 32760, 32760, variant)
    if "xattn" in which:  # cross-attention at the c3 launch shape: shared-S kernel (4) vs the short-key kernel (6)
        for variant in (4, 6, 7, 8, 9):
            bench_attn(2, 40, 32760, 512, variant)
        for variant in (6, 7, 9):   # c2's head count (12 heads) and a short prompt
            bench_attn(2, 12, 32760, 512, variant)
            bench_attn(2, 40, 32760, 192, variant)
    if "attn" in which:
        for variant in [int(v) for v in os.environ.get("VC_ATTN_VARIANTS", "0,1").split(",")]:
            bench_attn(1, 8, 32760, 32760, variant)
            bench_attn(2, 40, 32760, 512, variant)
