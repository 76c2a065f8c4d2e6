"""Run under torchrun on N GPUs: the Ulysses sequence-parallel forward must equal the single-GPU forward of the same
model and inputs (P-invariance).  Every rank holds the full weights, so each rank computes the
P=1 result locally and compares it with the gathered SP result."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from versecrafter_b200.transformer import VerseCrafterWanTransformer3DModel  # noqa: E402


def main():
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    cfg = dict(dim=1024, ffn_dim=2048, num_heads=8, num_layers=4, text_dim=256, geoada_in_dim=128)
    grid = (4, 32, 48)  # L = 1536, divisible by 8
    model = VerseCrafterWanTransformer3DModel(**cfg, device="cuda", dtype=torch.bfloat16, init=False).synthetic_init_(0)
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(2, 16, *grid, generator=g, device="cuda").bfloat16()
    geo = torch.randn(2, 128, *grid, generator=g, device="cuda").bfloat16()
    ctx = [torch.randn(n, 256, generator=g, device="cuda").bfloat16() * 0.1 for n in (37, 512)]
    t = torch.full((2,), 500.0, device="cuda")
    L = grid[0] * grid[1] * grid[2] // 4
    # second case: a token count that does NOT divide by the world size (and seq_len > L): the zero-padded tail tokens
    # must not be attended (k_lens), exactly as on one GPU
    grid_b = (3, 14, 26)   # L = 3*7*13 = 273
    xb = torch.randn(2, 16, *grid_b, generator=g, device="cuda").bfloat16()
    geob = torch.randn(2, 128, *grid_b, generator=g, device="cuda").bfloat16()
    Lb = 273 + 5
    cases = [("divisible", (x, t, geo, ctx, L)), ("ragged+padded", (xb, t, geob, ctx, Lb))]
    singles = [model(*a).float() for _, a in cases]
    model.enable_multi_gpus_inference()
    all_ok = True
    # NCCL all_to_all path, the peer-memory path fused into the kernels, then the same with the two-chain interleaved
    # schedule and split signal / wait barriers (transformer._run_blocks_interleaved)
    for mode in ("nccl", "fused", "fused+interleave"):
        os.environ["VC_SP_MODE"] = mode.split("+")[0]
        model.sp_interleave = "1" if mode.endswith("interleave") else "0"
        model._px, model._px_failed = None, False
        for (name, a), y1 in zip(cases, singles):
            yp = model(*a).float()
            yp2 = model(*a).float()  # second call: buffers are reused across steps
            torch.cuda.synchronize()
            if model._px is not None:
                model._px.check_status(wait=True)   # no barrier of these forwards timed out
            used = "fused" if model._px is not None else "nccl"
            cos = float(torch.nn.functional.cosine_similarity(y1.flatten().double(), yp.flatten().double(), dim=0))
            rel = float((y1 - yp).norm() / y1.norm())
            ok = cos > 0.9999 and rel < 1e-2 and torch.equal(yp, yp2) and used == mode.split("+")[0]
            if mode.endswith("interleave"):
                ok = ok and model._interleave_active() and torch.equal(yp, y1)
            flags = torch.tensor([1.0 if ok else 0.0], device="cuda")
            dist.all_reduce(flags, op=dist.ReduceOp.MIN)
            all_ok = all_ok and flags.item() == 1
            if rank == 0:
                print(f"sp_check world={world} mode={mode} (ran {used}) case={name}: cosine {cos:.7f} rel-L2 {rel:.5f} "
                      f"bit-identical-to-1gpu={torch.equal(yp, y1)} repeatable={torch.equal(yp, yp2)} -> "
                      f"{'OK' if flags.item() == 1 else 'FAIL'}", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if all_ok else 1)


if __name__ == "__main__":
    main()
