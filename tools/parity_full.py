"""north_star acceptance test at FULL depth and width (VERDICT r1 "next" #1): the whole denoise-step forward, CFG batch 2,
free-running, of the drop-in transformer against the oracle on the same GPU, same weights, same inputs.

    python tools/parity_full.py --config c2 [--fp32] [--out gpurun_out/parity_c2.json]
    python tools/parity_full.py --config c3 [--fp32]

Reports, per config:
  * final one-step latent cosine and rel-L2 of the candidate vs the bf16 oracle (`restate.forward(dtype=bf16)`, attention
    through flash_attn when importable — the reference's own GPU attention — else fp32 SDPA) — north_star: cosine >= 0.999;
  * with --fp32 the same against the fp32 oracle ("truth": bf16-valued weights, every op in fp32) and the ratio
    err_cand / err_ref16 (is the candidate farther from the truth than the reference's own bf16 arithmetic is?);
  * teacher-forced per-block rel-L2 (candidate block fed the oracle's recorded input) for a spread of backbone and
    GeoAdapter blocks — north_star: <= 1e-2.
Checker-side use of oracle/ (test infrastructure); the candidate path is versecrafter_b200 only.
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
from oracle import restate  # noqa: E402


def relerr(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def cosine(a, b):
    return float(torch.nn.functional.cosine_similarity(a.double().flatten(), b.double().flatten(), dim=0))


def run(config="c2", fp32=False, seed=0, attention="auto", blocks=None, geo_blocks=None, log=print):
    from versecrafter_b200.transformer import VerseCrafterWanTransformer3DModel
    cfg, grid, desc = CONFIGS[config]
    nb = cfg["num_layers"]
    ng = len(range(0, nb, 2))
    blocks = sorted({0, nb // 2 - 1, nb - 1}) if blocks is None else blocks
    geo_blocks = sorted({0, ng - 1}) if geo_blocks is None else geo_blocks
    torch.cuda.empty_cache()
    t0 = time.time()
    sd0 = restate.synth_state_dict(cfg, seed=seed, device="cuda", dtype=torch.bfloat16)
    model = VerseCrafterWanTransformer3DModel(**cfg, device="cuda", dtype=torch.bfloat16, init=False)
    model.load_state_dict(sd0, strict=True)
    del sd0
    model.eval()
    model.cache_step_invariants = False
    sd = dict(model.state_dict())          # the oracle reads the candidate's own storage: one copy of the weights
    inp = restate.synth_inputs(cfg, grid, seed=seed, device="cuda")
    args = (inp["x"], inp["t"], inp["geoada_context"], inp["context"], inp["seq_len"])
    log(f"[parity_full] {config}: weights + inputs ready in {time.time() - t0:.1f} s "
        f"({torch.cuda.memory_allocated() / 2**30:.1f} GiB allocated)")

    used = "sdpa"
    if attention in ("auto", "flash"):
        try:
            restate.set_attention_impl("flash")
            used = "flash_attn"
        except Exception as e:  # noqa: BLE001
            if attention == "flash":
                raise
            log(f"[parity_full] flash_attn unavailable ({e!r}); bf16 oracle attention = fp32 SDPA")
    keep = ["e0", "ctx"]
    for i in blocks:
        keep += [f"blk_in.{i}", f"blk_core.{i}"]
    for j in geo_blocks:
        keep += [f"geo_in.{j}", f"geo_out.{j}"]
    rec16 = restate.KeepKeys(keep)
    t0 = time.time()
    with torch.no_grad():
        y16 = restate.forward(sd, cfg, *args, dtype=torch.bfloat16, record=rec16)
    torch.cuda.synchronize()
    t_oracle16 = time.time() - t0
    restate.set_attention_impl("sdpa")
    log(f"[parity_full] bf16 oracle ({used}) forward: {t_oracle16:.1f} s")

    t0 = time.time()
    y = model(x=inp["x"].bfloat16(), t=inp["t"], geoada_context=inp["geoada_context"].bfloat16(), context=inp["context"],
              seq_len=inp["seq_len"], geoada_context_scale=1.0)
    torch.cuda.synchronize()
    t_cand = time.time() - t0
    out = {"config": config, "workload": desc, "blocks": nb, "geoada_blocks": ng, "batch": int(y.shape[0]),
           "tokens": inp["seq_len"], "seed": seed, "bf16_oracle_attention": used,
           "final_vs_bf16_oracle": {"cosine": cosine(y, y16), "rel_l2": relerr(y, y16)},
           "seconds": {"bf16_oracle": t_oracle16, "candidate_first_call": t_cand}}
    log(f"[parity_full] candidate vs bf16 oracle: cosine {out['final_vs_bf16_oracle']['cosine']:.6f} "
        f"rel-L2 {out['final_vs_bf16_oracle']['rel_l2']:.3e}")

    # teacher-forced per-block
    gridp = (grid[0], grid[1] // 2, grid[2] // 2)
    tf = {}
    for kind, idxs, kin, kout in (("blocks", blocks, "blk_in", "blk_core"), ("geoada_blocks", geo_blocks, "geo_in", "geo_out")):
        for i in idxs:
            got = model.block_forward(kind, i, rec16[f"{kin}.{i}"], rec16["e0"], rec16["ctx"], gridp)
            tf[f"{kind}.{i}"] = relerr(got, rec16[f"{kout}.{i}"])
            del got
    out["teacher_forced_block_rel_l2"] = tf
    out["teacher_forced_max"] = max(tf.values())
    log(f"[parity_full] teacher-forced per-block rel-L2: {json.dumps({k: round(v, 5) for k, v in tf.items()})}")
    rec16.clear()

    if fp32:
        t0 = time.time()
        with torch.no_grad():   # bf16-valued weights converted op by op (restate.linear casts w to x.dtype): no fp32 copy
            y32 = restate.forward(sd, cfg, *args, dtype=torch.float32)
        torch.cuda.synchronize()
        out["seconds"]["fp32_oracle"] = time.time() - t0
        e_c, e_r = relerr(y, y32), relerr(y16, y32)
        out["final_vs_fp32_oracle"] = {"cosine": cosine(y, y32), "rel_l2": e_c}
        out["bf16_oracle_vs_fp32_oracle"] = {"cosine": cosine(y16, y32), "rel_l2": e_r}
        out["err_cand_over_err_ref16"] = e_c / max(e_r, 1e-30)
        log(f"[parity_full] vs fp32 oracle: candidate cosine {out['final_vs_fp32_oracle']['cosine']:.6f} rel-L2 {e_c:.3e}; "
            f"bf16 oracle rel-L2 {e_r:.3e}; ratio {out['err_cand_over_err_ref16']:.3f} ({out['seconds']['fp32_oracle']:.1f} s)")
    out["pass"] = bool(out["final_vs_bf16_oracle"]["cosine"] >= 0.999 and out["teacher_forced_max"] <= 1e-2 and
                       (not fp32 or out["final_vs_fp32_oracle"]["cosine"] >= 0.999))
    model.release_workspace()
    del model, sd
    torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c2", choices=sorted(CONFIGS))
    ap.add_argument("--fp32", action="store_true")
    ap.add_argument("--attention", default="auto", choices=["auto", "flash", "sdpa"])
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    res = run(a.config, fp32=a.fp32, seed=a.seed, attention=a.attention)
    print(json.dumps(res), flush=True)
    if a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        with open(a.out, "w") as f:
            json.dump(res, f, indent=1)
    sys.exit(0 if res["pass"] else 1)


if __name__ == "__main__":
    main()
