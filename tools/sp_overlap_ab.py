"""Same-process A/B of the Ulysses exchange schedules at a bench config (run under torchrun on N >= 2 GPUs):
    one_scatter      fused QKV GEMM, then ONE scatter kernel (round-1 path, sp.ulysses_attention_fused)
    v_only           q|k GEMM + one scatter launch on the main stream, v scattered by its GEMM's epilogue
    full             q GEMM | k GEMM + scatter(q) on a side stream | v GEMM-scatter + scatter(k) on the side stream
each with per-thread stores or bulk (TMA-engine) stores, and with caps on the side-stream kernels' resident CTAs per SM.
The model is built once; every variant is warmed, timed over `--steps` forwards with CUDA events (max over ranks) and
checked bit-identical to the first variant's output.  One JSON line per variant on rank 0."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from versecrafter_b200 import ops  # noqa: E402
from versecrafter_b200.transformer import VerseCrafterWanTransformer3DModel  # noqa: E402

VARIANTS = [
    ("one_scatter_torch_barrier", dict(overlap=False, bulk=False, cap=0, barrier="torch")),
    ("one_scatter", dict(overlap=False, bulk=False, cap=0)),
    ("one_scatter_bulk", dict(overlap=False, bulk=True, cap=0)),
    ("v_only", dict(overlap=True, mode="v_only", bulk=False, cap=0)),
    ("v_only_bulk", dict(overlap=True, mode="v_only", bulk=True, cap=0)),
    ("full", dict(overlap=True, mode="full", bulk=False, cap=0)),
    ("full_cap2", dict(overlap=True, mode="full", bulk=False, cap=2)),
    ("full_bulk", dict(overlap=True, mode="full", bulk=True, cap=0)),
    ("full_bulk_cap1", dict(overlap=True, mode="full", bulk=True, cap=1)),
    ("full_bulk_cap2", dict(overlap=True, mode="full", bulk=True, cap=2)),
    ("interleave", dict(overlap=False, bulk=False, cap=0, interleave=True)),   # two-chain schedule, split signal / wait barriers
    ("one_scatter", dict(overlap=False, bulk=False, cap=0)),   # again: drift of the box over the run
    ("one_scatter_torch_barrier", dict(overlap=False, bulk=False, cap=0, barrier="torch")),
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c3")
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--only", default="")
    ap.add_argument("--seq", default="", help="comma-separated variant names, run in this order (repeats allowed)")
    args = ap.parse_args()
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    cfg, grid, desc = bench.CONFIGS[args.config]
    model = VerseCrafterWanTransformer3DModel(**cfg, device="cuda", dtype=torch.bfloat16, init=False)
    model.synthetic_init_(seed=0)
    model.cache_step_invariants = False
    model.eval()
    model.enable_multi_gpus_inference()
    host = bench.make_inputs(cfg, grid, 2)
    dev = dict(x=host["x"].cuda(), geo=host["geo"].cuda(), ctx=[c.cuda() for c in host["ctx"]], t=host["t"].cuda())

    def step():
        return model(x=dev["x"], t=dev["t"], geoada_context=dev["geo"], context=dev["ctx"], seq_len=host["seq_len"],
                     geoada_context_scale=1.0)

    first = None
    only = [s for s in args.only.split(",") if s]
    todo = VARIANTS
    if args.seq:
        table = dict(VARIANTS)
        todo = [(n, table[n]) for n in args.seq.split(",")]
    for name, v in todo:
        if only and name not in only:
            continue
        model.sp_interleave = "1" if v.get("interleave") else "0"
        px = model._peer_exchange()
        assert px is not None, "symmetric memory unavailable"
        model.sp_overlap = v["overlap"]
        px.overlap_mode = v.get("mode", "full")
        px.scatter_bulk, px.scatter_cap = v["bulk"], v["cap"]
        px.barrier_impl = v.get("barrier", "own")
        for _ in range(args.warmup):
            y = step()
        ops.timing = {}
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            y = step()
        e1.record()
        dist.barrier(); torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / args.steps], device="cuda")
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        per = {k: round(sum(a.elapsed_time(b) for a, b, _ in lst) / args.steps, 1) for k, lst in ops.timing.items()}
        ops.timing = None
        bar = torch.tensor([per.get("barrier", 0.0)], device="cuda")   # time inside the barriers: min over ranks = pure
        bmin, bmax = bar.clone(), bar.clone()                          # cost on the slowest rank, max = cost + rank skew
        dist.all_reduce(bmin, op=dist.ReduceOp.MIN)
        dist.all_reduce(bmax, op=dist.ReduceOp.MAX)
        per["barrier_min_rank"], per["barrier_max_rank"] = round(float(bmin), 2), round(float(bmax), 2)
        if first is None:
            first = y.clone()
        same = torch.tensor([1.0 if torch.equal(y, first) else 0.0], device="cuda")
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        if rank == 0:
            print(json.dumps({"variant": name, "config": args.config, "n_gpus": world, "ms_per_step": round(float(ms), 2),
                              "bit_identical_to_first": bool(same.item()), "kernel_ms_per_step": per}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
