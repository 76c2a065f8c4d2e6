"""BASELINE config c5: the full sampling loop (default 50 steps, UniPC, CFG) of a Wan2.2-A14B-style two-expert model
+ GeoAdapter at 81x480x832 on N GPUs (torchrun) or one GPU.  Synthetic weights / latents / control maps / prompts (no
checkpoints offline).  Device time by CUDA events around the whole loop, max over ranks; rank 0 prints one JSON line.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/c5_loop.py [--steps 50]
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from versecrafter_b200 import loop, ops, schedulers  # noqa: E402
from versecrafter_b200.expert_switch import VerseCrafterWan22ExpertSwitch  # noqa: E402
from versecrafter_b200.transformer import VerseCrafterWanTransformer3DModel  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--shift", type=float, default=12.0)          # config/wan2.2/wan_civitai_t2v.yaml scheduler_kwargs
    ap.add_argument("--guidance", type=float, default=6.0)
    ap.add_argument("--small", action="store_true", help="dim-1024 / 2-layer experts on a small grid (plumbing check)")
    ap.add_argument("--step-cache", type=int, default=1)
    args = ap.parse_args()
    rank, world, lrk = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(lrk)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", lrk))
    if args.small:
        cfg = dict(dim=1024, ffn_dim=2048, num_heads=8, num_layers=2, text_dim=256, geoada_in_dim=128)
        grid = (4, 32, 48)
    else:
        cfg = dict(dim=5120, ffn_dim=13824, num_heads=40, num_layers=40, text_dim=4096, geoada_in_dim=128)
        grid = (21, 60, 104)
    experts = [VerseCrafterWanTransformer3DModel(**cfg, device="cuda", dtype=torch.bfloat16, init=False).synthetic_init_(s)
               for s in (0, 1)]
    for m in experts:
        m.cache_step_invariants = bool(args.step_cache)
    model = VerseCrafterWan22ExpertSwitch(experts[0], experts[1], boundary=0.875).share_workspace().eval()
    if world > 1:
        model.enable_multi_gpus_inference()
    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.randn(1, 16, *grid, generator=g, device="cuda").bfloat16()
    lat = [torch.randn(16, *grid, generator=g, device="cuda").bfloat16() for _ in range(4)]
    mask = (torch.rand(1, 4 * (grid[0] - 1) + 1, 8 * grid[1], 8 * grid[2], generator=g, device="cuda") > 0.5).bfloat16()
    geo = loop.geoada_context(lat, mask)
    ctx = [torch.randn(n, cfg["text_dim"], generator=g, device="cuda").bfloat16() * 0.1 for n in (37, 512)]

    def run(n_steps):
        sch = schedulers.FlowUniPCMultistepScheduler(shift=1)
        sch.set_timesteps(n_steps, device="cuda", shift=args.shift)
        model.calls = {"high": 0, "low": 0}
        return loop.denoise(model, x, geo, ctx, guidance_scale=args.guidance, scheduler=sch), sch

    run(2)                                                   # warm-up: workspaces, symmetric-memory rendezvous
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    n0 = ops.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out, sch = run(args.steps)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    fin = torch.tensor([float(torch.isfinite(out.float()).all())], device="cuda")
    chk = out.float().sum().reshape(1)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(fin, op=dist.ReduceOp.MIN)
        lo, hi = chk.clone(), chk.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        same = bool(lo.item() == hi.item())
    else:
        same = True
    L = grid[0] * (grid[1] // 2) * (grid[2] // 2)
    if rank == 0:
        s = ms.item() / 1e3
        print(json.dumps({
            "workload": "c5: two-expert (boundary 0.875) 14B+GeoAdapter, 81x480x832, CFG 2, UniPC" if not args.small else "c5-small",
            "n_gpus": world, "steps": args.steps, "shift": args.shift, "loop_s": s, "s_per_step": s / args.steps,
            "video_tokens_per_s": 2 * L * args.steps / s, "expert_calls": model.calls,
            "timesteps_head": sch.timesteps[:4].tolist(), "timesteps_tail": sch.timesteps[-3:].tolist(),
            "finite": bool(fin.item()), "ranks_agree": same, "step_cache": bool(args.step_cache),
            "gpu_launches": ops.launch_count - n0, "out_abs_mean": float(out.float().abs().mean()),
            "data": "synthetic", "mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30}), flush=True)
    if world > 1:
        dist.destroy_process_group()
    sys.exit(0 if (fin.item() == 1 and same) else 1)


if __name__ == "__main__":
    main()
