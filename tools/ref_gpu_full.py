"""The reference's GPU path at FULL size on the same B200, as parity target and as timed baseline (VERDICT r2 "next" #1;
BASELINE.md §4 row B-GPU-ref): the reference's own `VerseCrafterWanTransformer3DModel`
(wan_transformer3d_versecrafter.py:294-442, imported unmodified through oracle/ref_import.py from oracle/_ref) run the way
the pipeline runs it (pipeline_wan_versecrafter.py:893-901) — bf16 weights, torch.autocast("cuda", bf16), attention() bound
to flash_attn — next to the drop-in, same weights, same inputs, whole denoise-step forward with CFG batch 2.

    python tools/ref_gpu_full.py --config c2 [--iters 3] [--out gpurun_out/ref_gpu_full_c2.json]
    python tools/ref_gpu_full.py --config c3

Reports ref_ms / repo_ms (CUDA events, 1 warm-up + `iters` forwards each), speedup, final cosine / rel-L2 of the drop-in
against the reference's output, and teacher-forced rel-L2 of blocks 0 / mid / last and GeoAdapter blocks 0 / last (the
drop-in's block fed the input the reference's block saw, captured with forward hooks; for GeoAdapter blocks the stacked
output's [-1] / [-2] are the block output and the hint).  The reference model is built on the meta device and takes the
drop-in's parameter storage by `load_state_dict(assign=True)`: ONE copy of the weights serves both, which is what lets
the 14B config fit next to the reference's 14 GB hint stack.  Checker-side tooling; nothing here is on the product path.
"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from bench import CONFIGS  # noqa: E402
from oracle import ref_import, restate  # noqa: E402


def relerr(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def cosine(a, b):
    return float(torch.nn.functional.cosine_similarity(a.double().flatten(), b.double().flatten(), dim=0))


def reference_model_sharing(model, cfg, device="cuda"):
    """The reference's model class carrying the drop-in's parameters WITHOUT a second copy: constructed on the meta device
    (no 14B random init), parameters assigned from the drop-in's state dict (same names, strict), `freqs` (a plain
    attribute there, wan_transformer3d.py:775-783) rebuilt with the reference's own rope_params."""
    base, vc = ref_import.load_reference_modules()
    with torch.device("meta"):
        ref = vc.VerseCrafterWanTransformer3DModel(**cfg)
    sd = {k: v for k, v in model.state_dict().items()}
    ref.load_state_dict(sd, strict=True, assign=True)
    d = cfg["dim"] // cfg["num_heads"]
    ref.freqs = torch.cat([base.rope_params(1024, d - 4 * (d // 6)), base.rope_params(1024, 2 * (d // 6)),
                           base.rope_params(1024, 2 * (d // 6))], dim=1)
    leftovers = [n for n, p in list(ref.named_parameters()) + list(ref.named_buffers()) if p.is_meta]
    assert not leftovers, f"still on meta: {leftovers[:5]}"
    return ref.eval().requires_grad_(False)


def _timed(fn, iters):
    fn()                                     # warm-up
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
def run(config="c2", iters=3, seed=0, blocks=None, geo_blocks=None, timing=True, log=print):
    from versecrafter_b200.transformer import VerseCrafterWanTransformer3DModel
    base, _ = ref_import.load_reference_modules()      # also puts the stand-in videox_fun package on sys.path
    from videox_fun.models import attention_utils
    cfg, grid, desc = CONFIGS[config]
    nb = cfg["num_layers"]
    ng = len(range(0, nb, 2))
    blocks = sorted({0, nb // 2 - 1, nb - 1}) if blocks is None else blocks
    geo_blocks = sorted({0, ng - 1}) if geo_blocks is None else geo_blocks
    gridp = (grid[0], grid[1] // 2, grid[2] // 2)
    torch.cuda.empty_cache()
    t0 = time.time()
    model = VerseCrafterWanTransformer3DModel(**cfg, device="cuda", dtype=torch.bfloat16, init=False)
    model.synthetic_init_(seed=seed)
    model.eval()
    model.cache_step_invariants = False
    ref = reference_model_sharing(model, cfg)
    inp = restate.synth_inputs(cfg, grid, seed=seed, device="cuda")
    x, geo = inp["x"].bfloat16(), inp["geoada_context"].bfloat16()
    ctx = [c.bfloat16() for c in inp["context"]]
    log(f"[ref_gpu_full] {config}: one copy of the weights + inputs ready in {time.time() - t0:.1f} s "
        f"({torch.cuda.memory_allocated() / 2**30:.1f} GiB allocated)")

    # -- the reference's forward, hooks on the selected blocks only (a clone of every block's input would not fit at c3) --
    cap, handles = {}, []

    def backbone_hooks(i):
        def pre(mod, args, kwargs):
            cap[f"blk_in.{i}"] = args[0].detach().clone()
            cap["e0"], cap["ctx"] = kwargs["e"], kwargs["context"]
            cap[f"blk_kwargs.{i}"] = {k: v for k, v in kwargs.items() if k not in ("hints", "context_scale")}
        return pre

    def geo_hooks(j):
        def pre(mod, args, kwargs):
            c = args[0]
            cap[f"geo_in.{j}"] = (mod.before_proj(c) + kwargs["x"]).detach().clone() if j == 0 else c[-1].detach().clone()
        def post(mod, args, kwargs, out):
            cap[f"geo_out.{j}"] = out[-1].detach().clone()
        return pre, post

    for i in blocks:
        handles.append(ref.blocks[i].register_forward_pre_hook(backbone_hooks(i), with_kwargs=True))
    for j in geo_blocks:
        pre, post = geo_hooks(j)
        handles += [ref.geoada_blocks[j].register_forward_pre_hook(pre, with_kwargs=True),
                    ref.geoada_blocks[j].register_forward_hook(post, with_kwargs=True)]

    def ref_forward():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            return ref(x=x, t=inp["t"], geoada_context=geo, context=ctx, seq_len=inp["seq_len"], geoada_context_scale=1.0)

    torch.cuda.reset_peak_memory_stats()
    y_ref = ref_forward()
    assert attention_utils.LAST_IMPL == "flash_attn", attention_utils.LAST_IMPL
    with torch.autocast("cuda", dtype=torch.bfloat16):   # block cores by the reference's own WanAttentionBlock.forward
        for i in blocks:
            cap[f"blk_core.{i}"] = base.WanAttentionBlock.forward(ref.blocks[i], cap[f"blk_in.{i}"], **cap.pop(f"blk_kwargs.{i}"))
    for h in handles:
        h.remove()
    ref_peak = torch.cuda.max_memory_allocated() / 2**30
    ref_ms = ref_min = None
    if timing:
        ref_ms, ref_min = _timed(ref_forward, iters)
    log(f"[ref_gpu_full] reference forward: {ref_ms} ms (peak {ref_peak:.1f} GiB)")
    torch.cuda.empty_cache()

    # -- the drop-in on the same inputs -------------------------------------------------------------------------------
    def repo_forward():
        return model(x=x, t=inp["t"], geoada_context=geo, context=ctx, seq_len=inp["seq_len"], geoada_context_scale=1.0)

    y = repo_forward()
    final = dict(cosine=cosine(y, y_ref), rel_l2=relerr(y, y_ref))
    repo_ms = repo_min = None
    if timing:
        repo_ms, repo_min = _timed(repo_forward, iters)
    tf = {}
    for i in blocks:
        got = model.block_forward("blocks", i, cap[f"blk_in.{i}"], cap["e0"], cap["ctx"], gridp)
        tf[f"blocks.{i}"] = relerr(got, cap[f"blk_core.{i}"])
    for j in geo_blocks:
        got = model.block_forward("geoada_blocks", j, cap[f"geo_in.{j}"], cap["e0"], cap["ctx"], gridp)
        tf[f"geoada_blocks.{j}"] = relerr(got, cap[f"geo_out.{j}"])
    L = inp["seq_len"]
    rec = dict(what="whole denoise-step forward (CFG batch 2): the reference's own VerseCrafterWanTransformer3DModel under "
                    "autocast(bf16) + flash_attn vs this repo's drop-in, same weights (shared storage) and inputs, 1xB200",
               config=config, workload=desc, tokens_per_step=2 * L, ref_ms=ref_ms, ref_ms_min=ref_min, repo_ms=repo_ms,
               repo_ms_min=repo_min, speedup=(ref_ms / repo_ms) if timing else None,
               ref_tokens_per_s=(2 * L / (ref_ms / 1e3)) if timing else None,
               repo_tokens_per_s=(2 * L / (repo_ms / 1e3)) if timing else None,
               final_cosine=final["cosine"], final_rel_l2=final["rel_l2"], teacher_forced_rel_l2=tf,
               worst_teacher_forced=max(tf.values()), iters=iters, reference_attention=attention_utils.LAST_IMPL,
               reference_source=os.path.relpath(ref_import.REF_ROOT, ROOT) if ref_import.REF_ROOT.startswith(ROOT) else ref_import.REF_ROOT,
               reference_peak_gib=ref_peak, gpu=torch.cuda.get_device_name(0))
    del ref, model, cap
    torch.cuda.empty_cache()
    return rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c2", choices=sorted(CONFIGS))
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    if not ref_import.reference_available():
        raise SystemExit("reference sources absent: run `python -m oracle.build_ref` where /root/reference exists")
    rec = run(a.config, a.iters)
    print(json.dumps(rec), flush=True)
    out = a.out or os.path.join(ROOT, "gpurun_out", f"ref_gpu_full_{a.config}.json")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    with open(out, "w") as f:
        f.write(json.dumps(rec) + "\n")


if __name__ == "__main__":
    main()
