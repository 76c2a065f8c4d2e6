"""Diagnostic (run under torchrun on >= 2 GPUs): first per-block tensor at which the sequence-parallel forward leaves the
single-GPU forward.  Same model / inputs as tools/sp_check.py's "divisible" case; SPD_LAYERS, SPD_DIM, SPD_HEADS override."""
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
    nl = int(os.environ.get("SPD_LAYERS", "4"))
    dim, heads = int(os.environ.get("SPD_DIM", "1024")), int(os.environ.get("SPD_HEADS", "8"))
    cfg = dict(dim=dim, ffn_dim=2 * dim, num_heads=heads, num_layers=nl, text_dim=256, geoada_in_dim=128)
    grid = (4, 32, 48)
    model = VerseCrafterWanTransformer3DModel(**cfg, device="cuda", dtype=torch.bfloat16, init=False).synthetic_init_(0)
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(2, 16, *grid, generator=g, device="cuda").bfloat16()
    geo = torch.randn(2, 128, *grid, generator=g, device="cuda").bfloat16()
    ctx = [torch.randn(n, 256, generator=g, device="cuda").bfloat16() * 0.1 for n in (37, 512)]
    t = torch.full((2,), 500.0, device="cuda")
    L = grid[0] * grid[1] * grid[2] // 4
    model.cache_step_invariants = False
    model.record = {}
    y1 = model(x, t, geo, ctx, L)
    rec1, model.record = model.record, None
    model.enable_multi_gpus_inference()
    for mode in os.environ.get("SPD_MODES", "nccl,fused").split(","):
        os.environ["VC_SP_MODE"] = mode
        model._px, model._px_failed = None, False
        model.record = {}
        yp = model(x, t, geo, ctx, L)
        recp, model.record = model.record, None
        torch.cuda.synchronize()
        Lp = L // world
        for r in range(world):
            dist.barrier()
            if r != rank:
                continue
            first = None
            for k, v in recp.items():
                ref = rec1[k]
                if v.dim() == 3 and ref.shape[1] == L and v.shape[1] == Lp:
                    ref = ref[:, rank * Lp:(rank + 1) * Lp]
                if ref.shape != v.shape:
                    continue
                nd = int((ref != v).sum())
                if nd and first is None:
                    first = k
                    rows = (ref != v).any(dim=-1).nonzero()
                    cols = (ref != v).any(dim=0).any(dim=0).nonzero().flatten() if v.dim() == 3 else None
                    print(f"[{mode} layers={nl} rank {rank}] FIRST difference at {k}: {nd} of {v.numel()} elements, max abs "
                          f"{float((ref.float() - v.float()).abs().max()):.4g}; rows(b,t) first {rows[:4].tolist()} last "
                          f"{rows[-2:].tolist()} n_rows {rows.shape[0]}; cols {cols[:3].tolist() if cols is not None else ''}..."
                          f"{cols[-3:].tolist() if cols is not None else ''} n_cols {cols.numel() if cols is not None else ''}",
                          flush=True)
            print(f"[{mode} layers={nl} rank {rank}] final bit-identical={torch.equal(yp, y1)} first_diff={first}", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
