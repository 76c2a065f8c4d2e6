"""The bar on the same B200: one DiT block of the reference's op sequence in eager PyTorch (separate q/k/v Linears on
cuBLASLt, fp32 LayerNorm + unfused modulate, fp32 RMSNorm, complex fp64 RoPE, flash-attn 2.8 `flash_attn_func`, unfused
GELU-tanh FFN; SURVEY.md App. A.3-A.7, wan_transformer3d.py:564-611) timed at the c3 launch shape
(B=2, L=32,760, D=5120, F=13824, H=40) next to the same block on this repo's kernels.  Library code only on the
reference side; not on the product path.  Writes one JSON line to gpurun_out/ref_gpu_block.json.

    python tools/ref_gpu_block.py [--iters 3]
"""
import argparse
import json
import math
import os
import sys
from types import SimpleNamespace

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from versecrafter_b200.transformer import VerseCrafterWanTransformer3DModel  # noqa: E402


def rms(u, g, eps=1e-6):
    return (u.float() * torch.rsqrt(u.float().pow(2).mean(-1, keepdim=True) + eps)).type_as(u) * g


def rope_eager(u, grid, freqs):
    """[B, L, H, 128] -> rotated; complex128 tables split 22|21|21 over (frame, height, width), as the reference."""
    B, L, H, d = u.shape
    f, h, w = grid
    c = d // 2
    parts = freqs.split([c - 2 * (c // 3), c // 3, c // 3], dim=1)
    tab = torch.cat([parts[0][:f].view(f, 1, 1, -1).expand(f, h, w, -1), parts[1][:h].view(1, h, 1, -1).expand(f, h, w, -1),
                     parts[2][:w].view(1, 1, w, -1).expand(f, h, w, -1)], dim=-1).reshape(L, 1, -1)
    out = []
    for b in range(B):                                       # the reference loops over the batch
        z = torch.view_as_complex(u[b].to(torch.float64).reshape(L, H, -1, 2))
        out.append(torch.view_as_real(z * tab).flatten(2))
    return torch.stack(out).type_as(u)


def eager_block(x, m, ctx, blk, grid, freqs, eps=1e-6):
    from flash_attn import flash_attn_func
    B, L, D = x.shape
    H = blk.num_heads
    sa, ca = blk.self_attn, blk.cross_attn
    ln = lambda t: F.layer_norm(t.float(), (D,), eps=eps).type_as(t)
    lin = lambda t, mod: F.linear(t, mod.weight, mod.bias)
    e = [m[:, i:i + 1] for i in range(6)]
    h = ln(x) * (1 + e[1]) + e[0]
    q = rms(lin(h, sa.q), sa.norm_q.weight).view(B, L, H, -1)
    k = rms(lin(h, sa.k), sa.norm_k.weight).view(B, L, H, -1)
    v = lin(h, sa.v).view(B, L, H, -1)
    a = flash_attn_func(rope_eager(q, grid, freqs), rope_eager(k, grid, freqs), v).flatten(2)
    x = x + lin(a, sa.o) * e[2]
    h = F.layer_norm(x.float(), (D,), blk.norm3.weight.float(), blk.norm3.bias.float(), eps).type_as(x)
    q = rms(lin(h, ca.q), ca.norm_q.weight).view(B, L, H, -1)
    k = rms(lin(ctx, ca.k), ca.norm_k.weight).view(B, -1, H, 128)
    v = lin(ctx, ca.v).view(B, -1, H, 128)
    x = x + lin(flash_attn_func(q, k, v).flatten(2), ca.o)
    h = ln(x) * (1 + e[4]) + e[3]
    y = lin(F.gelu(lin(h, blk.ffn[0]), approximate="tanh"), blk.ffn[2])
    return x + y * e[5]


def timed(fn, iters):
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


@torch.no_grad()
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=3)
    args = ap.parse_args()
    cfg = dict(dim=5120, ffn_dim=13824, num_heads=40, num_layers=1, text_dim=4096, geoada_in_dim=128)
    grid = (21, 30, 52)
    L, B, D = grid[0] * grid[1] * grid[2], 2, cfg["dim"]
    model = VerseCrafterWanTransformer3DModel(**cfg, device="cuda", dtype=torch.bfloat16, init=False).synthetic_init_(0)
    blk = model.blocks[0]
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn(B, L, D, generator=g, device="cuda").bfloat16()
    e0 = torch.randn(B, 6, D, generator=g, device="cuda").bfloat16() * 0.1
    ctx = torch.randn(B, 512, D, generator=g, device="cuda").bfloat16()
    m = (blk.modulation + e0).to(torch.bfloat16)
    freqs = model.freqs.cuda()

    y_ref = eager_block(x, m, ctx, blk, grid, freqs)
    y = model.block_forward("blocks", 0, x, e0, ctx, grid)
    rel = float((y.float() - y_ref.float()).norm() / y_ref.float().norm())
    ref_ms, ref_min = timed(lambda: eager_block(x, m, ctx, blk, grid, freqs), args.iters)
    st = SimpleNamespace(B=B, Lp=L, grid=grid, rope=model._rope(grid), k_len=L, step_cache=None)
    x2d = x.reshape(B * L, D).clone()
    c2d = ctx.reshape(-1, D).contiguous()
    ours_ms, ours_min = timed(lambda: model._block(blk, x2d, m, c2d, st), args.iters * 2)
    flop = B * (12 * L * D * D + 4 * 512 * D * D + 4 * L * D * cfg["ffn_dim"] + 4 * L * L * D + 4 * L * 512 * D)
    rec = dict(what="one DiT block at the c3 shape, same weights and inputs, 1xB200", shape=[B, L, D], block_tflop=flop / 1e12,
               eager_pytorch_fa2_ms=ref_ms, eager_pytorch_fa2_ms_min=ref_min, eager_tflops=flop / ref_ms / 1e9,
               this_repo_ms=ours_ms, this_repo_ms_min=ours_min, this_repo_tflops=flop / ours_ms / 1e9,
               speedup=ref_ms / ours_ms, rel_l2_between_them=rel, step_extrapolation_s=dict(eager=ref_ms * 60 / 1e3, ours=ours_ms * 60 / 1e3),
               note="eager side excludes the reference's hint stack/unbind copies (~150 GB/step) and runs RoPE in fp64 as the reference does")
    print(json.dumps(rec), flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/ref_gpu_block.json", "w") as f:
        f.write(json.dumps(rec) + "\n")


if __name__ == "__main__":
    main()
