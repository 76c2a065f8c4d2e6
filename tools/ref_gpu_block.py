"""The bar on the same B200: one DiT block of the REFERENCE'S OWN `WanAttentionBlock` (imported unmodified through
oracle/ref_import.py from /root/reference or its byte-for-byte copy oracle/_ref; wan_transformer3d.py:524-611) run the way
the pipeline runs it — bf16 weights, torch.autocast("cuda", bf16), attention() bound to flash_attn 2.8 `flash_attn_func` —
timed at the c3 launch shape (B=2, L=32,760, D=5120, F=13824, H=40) next to the same block, same weights, on this repo's
kernels.  Measurement tooling (checker side), not on the product path.  Writes gpurun_out/ref_gpu_block.json.

    python tools/ref_gpu_block.py [--iters 3]
"""
import argparse
import json
import os
import sys
from types import SimpleNamespace

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_import  # noqa: E402
from versecrafter_b200.transformer import VerseCrafterWanTransformer3DModel  # noqa: E402


def reference_block_from(blk, cfg):
    """The reference's WanAttentionBlock carrying the drop-in block's weights (identical parameter names)."""
    base, _ = ref_import.load_reference_modules()
    ref = base.WanAttentionBlock("t2v_cross_attn", cfg["dim"], cfg["ffn_dim"], cfg["num_heads"], (-1, -1), True, True, 1e-6)
    missing, unexpected = ref.load_state_dict({k: v for k, v in blk.state_dict().items()}, strict=True)
    return ref.to(device="cuda", dtype=torch.bfloat16).eval().requires_grad_(False)


def reference_block_forward(ref, x, e0, ctx, grid, freqs):
    B, L, _ = x.shape
    with torch.autocast("cuda", dtype=torch.bfloat16):
        return ref(x, e0, torch.full((B,), L, dtype=torch.long), torch.tensor([list(grid)] * B), freqs, ctx, None,
                   dtype=torch.bfloat16)


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
def measure_block(model, cfg, grid, iters=3, B=2):
    """One DiT block (model.blocks[0]) at `grid` (latent F, H', W' AFTER patching), CFG batch B: the reference's module
    under autocast(bf16) + flash_attn vs the drop-in's kernels, same weights, same inputs.  Used by main() below and by
    bench.py's `gpu_reference` object (outside its timed region)."""
    L, D = grid[0] * grid[1] * grid[2], cfg["dim"]
    blk = model.blocks[0]
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn(B, L, D, generator=g, device="cuda").bfloat16()
    e0 = torch.randn(B, 6, D, generator=g, device="cuda").bfloat16() * 0.1
    ctx = torch.randn(B, 512, D, generator=g, device="cuda").bfloat16()
    m = (blk.modulation + e0).to(torch.bfloat16)
    freqs = model.freqs.cuda()
    ref = reference_block_from(blk, cfg)
    from videox_fun.models import attention_utils

    y_ref = reference_block_forward(ref, x, e0, ctx, grid, freqs)
    assert attention_utils.LAST_IMPL == "flash_attn", attention_utils.LAST_IMPL
    y = model.block_forward("blocks", 0, x, e0, ctx, grid)
    rel = float((y.float() - y_ref.float()).norm() / y_ref.float().norm())
    del y, y_ref
    ref_ms, ref_min = timed(lambda: reference_block_forward(ref, x, e0, ctx, grid, freqs), iters)
    st = SimpleNamespace(B=B, Lp=L, grid=grid, rope=model._rope(grid), k_len=L, l_true=L, step_cache=None)
    x2d = x.reshape(B * L, D).clone()
    c2d = ctx.reshape(-1, D).contiguous()
    ours_ms, ours_min = timed(lambda: model._block(blk, x2d, m, c2d, st), iters * 2)
    flop = B * (12 * L * D * D + 4 * 512 * D * D + 4 * L * D * cfg["ffn_dim"] + 4 * L * L * D + 4 * L * 512 * D)
    n_blocks = cfg["num_layers"] + len(range(0, cfg["num_layers"], 2))
    rec = dict(what="one DiT block: the reference's own WanAttentionBlock (eager PyTorch + cuBLASLt + flash-attn, autocast bf16) "
                    "vs this repo's kernels, same weights and inputs, same GPU",
               shape=[B, L, D], block_tflop=flop / 1e12, block_ms_ref=ref_ms, block_ms_ref_min=ref_min,
               ref_tflops=flop / ref_ms / 1e9, block_ms_repo=ours_ms, block_ms_repo_min=ours_min,
               repo_tflops=flop / ours_ms / 1e9, speedup=ref_ms / ours_ms, rel_l2_between_them=rel,
               reference_source=os.path.relpath(ref_import.REF_ROOT), reference_attention=attention_utils.LAST_IMPL,
               note="reference side excludes its hint stack / unbind copies and per-step embeddings (generous to it)")
    del ref, x, x2d, ctx, c2d
    torch.cuda.empty_cache()
    return rec


@torch.no_grad()
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=3)
    args = ap.parse_args()
    cfg = dict(dim=5120, ffn_dim=13824, num_heads=40, num_layers=1, text_dim=4096, geoada_in_dim=128)
    model = VerseCrafterWanTransformer3DModel(**cfg, device="cuda", dtype=torch.bfloat16, init=False).synthetic_init_(0)
    rec = measure_block(model, cfg, (21, 30, 52), args.iters)
    print(json.dumps(rec), flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/ref_gpu_block.json", "w") as f:
        f.write(json.dumps(rec) + "\n")


if __name__ == "__main__":
    main()
