"""bench.py — one denoise-step forward of Wan2.1-14B + GeoAdapter (BASELINE.json metric), B200-native path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config c3|c2|c1] [--impl b200|reference]
  python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...      (N > 1: Ulysses sequence parallel)

A "step" is one transformer forward (CFG batch 2, TeaCache off, cfg-skip off) on synthetic latents / control maps /
text embeddings of the named shape with random non-trivial bf16 weights.  `value` times K steps with the inputs
resident in HBM (CUDA events, max over ranks); `e2e` times the same K steps through the public forward() with the
inputs in pinned host memory (H2D copies and the D2H read of the prediction inside the timed region).
`--impl reference` times the oracle's CPU port of the reference path on the host cores (a bounded sample).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    # name: (ctor kwargs, latent (F, H, W), description)
    "c3": (dict(dim=5120, ffn_dim=13824, num_heads=40, num_layers=40, text_dim=4096, geoada_in_dim=128),
           (21, 60, 104), "Wan2.1-14B+GeoAdapter 81x480x832 (L=32760, CFG batch 2)"),
    "c4": (dict(dim=5120, ffn_dim=13824, num_heads=40, num_layers=40, text_dim=4096, geoada_in_dim=128),
           (21, 90, 160), "Wan2.1-14B+GeoAdapter 81x720x1280 (L=75600, CFG batch 2)"),
    "c2": (dict(dim=1536, ffn_dim=8960, num_heads=12, num_layers=30, text_dim=4096, geoada_in_dim=128),
           (21, 60, 104), "Wan2.1-1.3B-arch+GeoAdapter 81x480x832 (L=32760, CFG batch 2)"),
    "c1": (dict(dim=512, ffn_dim=2048, num_heads=4, num_layers=2, text_dim=4096, geoada_in_dim=128),
           (5, 30, 52), "reduced-depth DiT (2 blocks, dim 512)+GeoAdapter 5x30x52 (L=1950, CFG batch 2)"),
}


def load_peaks():
    p = dict(hbm_gbs=6650.0, bf16_tflops=1590.0, bf16_tflops_sustained=1400.0, source="fallback")
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p.update(json.load(f))
        p["source"] = "measured"
    except Exception:  # noqa: BLE001
        pass
    return p


def work_per_step(cfg, L, B, T=512):
    """Algorithmic FLOPs of one step."""
    D, Fd, nb = cfg["dim"], cfg["ffn_dim"], cfg["num_layers"]
    ng = len(cfg.get("geoada_layers") or range(0, nb, 2))
    gemm_blk = 8 * L * D * D + 4 * L * D * D + 4 * T * D * D + 4 * L * D * Fd
    attn_blk = 4 * L * L * D + 4 * L * T * D
    gemm = B * ((nb + ng) * gemm_blk + ng * 2 * L * D * D + 2 * L * D * D + 2 * L * D * (64 + 512 + 64))
    attn = B * (nb + ng) * attn_blk
    return float(gemm), float(attn)


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.limit")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "200"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons, pw, plim = [], None, set(), [], None
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
            except Exception:  # noqa: BLE001
                continue
            try:
                pw.append(float(r[2])); plim = float(r[7])
            except Exception:  # noqa: BLE001
                pass
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        pw.sort()
        # power_w_median / power_limit_w: the step runs AT the board power limit (reason sw_power_cap), which is why
        # schedule changes that remove idle time do not shorten it
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "power_w_median": pw[len(pw) // 2] if pw else None, "power_limit_w": plim}


def make_inputs(cfg, grid, B=2):
    """Synthetic inputs in pinned host memory, bf16."""
    g = torch.Generator().manual_seed(77)
    Fr, Hh, Ww = grid
    x = torch.randn(B, cfg.get("in_dim", 16), Fr, Hh, Ww, generator=g).to(torch.bfloat16)
    half = cfg["geoada_in_dim"] // 2
    ctrl = torch.cat([torch.randn(half, Fr, Hh, Ww, generator=g),
                      (torch.rand(cfg["geoada_in_dim"] - half, Fr, Hh, Ww, generator=g) > 0.5).float()]).to(torch.bfloat16)
    geo = torch.stack([ctrl] * B)
    ctx = [(0.1 * torch.randn(n, cfg["text_dim"], generator=g)).to(torch.bfloat16) for n in ([37, 512] * B)[:B]]
    t = torch.full((B,), 500.0)
    pin = lambda u: u.pin_memory() if torch.cuda.is_available() else u
    return dict(x=pin(x), geo=pin(geo), ctx=[pin(c) for c in ctx], t=pin(t), seq_len=Fr * (Hh // 2) * (Ww // 2))


# ------------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle's port of the reference path, timed on the host cores on a bounded sample
# ------------------------------------------------------------------------------------------------------------------
def _block_flops(cfg, Ls, T=512):
    D, Fd = cfg["dim"], cfg["ffn_dim"]
    return 12.0 * Ls * D * D + 4.0 * T * D * D + 4.0 * Ls * D * Fd + 4.0 * Ls * Ls * D + 4.0 * Ls * T * D


def bench_config(name, world):
    """The `config` object of the JSON line (shared by both arms so the driver sees the same workload keys)."""
    cfg, grid, desc = CONFIGS[name]
    L = grid[0] * (grid[1] // 2) * (grid[2] // 2)
    return {"workload": desc, "name": name, "tokens_per_step": 2 * L, "batch": 2, "seq_len": L,
            "parallelism": f"ulysses_sp{world}", "l2": "working set per kernel >> 126 MB L2 (no flush needed)",
            "teacache": "off", "cfg_skip": "off", "step_cache": "off", "weights": "random non-trivial bf16 (synthetic_init_)"}


def cpu_reference_sample(cfg_name, steps=2, total_budget=90.0, max_step_seconds=28.0):
    """CPU arm.  One "step" = one DiT block at the config's real width, fp32, on all host cores — the reference's OWN
    `WanAttentionBlock.forward` (wan_transformer3d.py:564-611), imported unmodified from /root/reference or its byte-for-byte
    copy under oracle/_ref (kind "reference"; attention() = torch SDPA, VideoX-Fun being un-vendored), else the oracle's
    restatement (kind "port").  The sample does NOT depend on --steps / --warmup: it is the config's FULL token grid whenever
    a 1-frame calibration says one full block fits max_step_seconds (c3 on 16 cores: 13-19 s), else the first `frames`
    latent frames of the real H'xW' grid that do.  Only the NUMBER of timed blocks follows `steps`, capped by total_budget
    (at least one; the calibration runs are the warm-up), so every caller on one box measures the same thing.
    Extrapolation to the metric (stated in `sample`): seconds per FULL-grid block = measured s/step x flops(full grid) /
    flops(sample) (x1 when the full grid ran; attention is quadratic in the token count); a denoise step = that x
    (num_layers + geoada blocks) x CFG batch 2; tokens/s = 2 L / s_per_denoise_step.  Embeddings, adapter projections and
    the hint copies are left out (generous to the CPU).  Returns (cpu_baseline dict, mean measured seconds per step,
    extrapolated s per denoise step, number of timed steps)."""
    from oracle import ref_import, restate
    cfg, grid, _ = CONFIGS[cfg_name]
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    D, H = cfg["dim"], cfg["num_heads"]
    gh, gw = grid[1] // 2, grid[2] // 2
    g = torch.Generator().manual_seed(0)
    e0 = torch.randn(1, 6, D, generator=g) * 0.1
    ctx = torch.randn(1, 512, D, generator=g)
    if ref_import.reference_available():
        kind = "reference"
        base, _vc = ref_import.load_reference_modules()
        torch.manual_seed(0)
        blk = base.WanAttentionBlock("t2v_cross_attn", D, cfg["ffn_dim"], H, (-1, -1), True, True, 1e-6).eval().requires_grad_(False)
        d = D // H
        freqs = torch.cat([base.rope_params(1024, d - 4 * (d // 6)), base.rope_params(1024, 2 * (d // 6)),
                           base.rope_params(1024, 2 * (d // 6))], dim=1)
        src = os.path.relpath(ref_import.REF_ROOT, ROOT) if ref_import.REF_ROOT.startswith(ROOT) else ref_import.REF_ROOT
        what = f"the reference's own WanAttentionBlock.forward ({src}, unmodified; attention() = torch SDPA)"

        def block(x, sgrid):
            return blk(x, e0, torch.tensor([x.shape[1]]), torch.tensor([list(sgrid)]), freqs, ctx, None, dtype=torch.float32)
    else:
        kind = "port"
        shapes = restate.block_param_shapes(D, cfg["ffn_dim"])
        sd = {"b." + k: (torch.randn(s, generator=g) * (0.02 if len(s) > 1 else 0.1) + (1.0 if "norm" in k and k.endswith("weight") else 0.0))
              for k, s in shapes.items()}
        what = "oracle/restate.dit_block (restatement of WanAttentionBlock.forward)"

        def block(x, sgrid):
            return restate.dit_block(sd, "b.", x, e0, restate.rope_angles(D // H, sgrid), ctx, H)

    def run_block(frames):
        sgrid = (frames, gh, gw)
        x = torch.randn(1, frames * gh * gw, D, generator=g)
        t0 = time.perf_counter()
        with torch.no_grad():
            block(x, sgrid)
        return time.perf_counter() - t0

    run_block(1)                                  # page in / thread pool
    t_cal = run_block(1)
    rate = _block_flops(cfg, gh * gw) / t_cal     # flop/s of this box on the 1-frame block
    step_budget = max_step_seconds
    frames = 1
    for f in range(grid[0], 0, -1):
        if _block_flops(cfg, f * gh * gw) / rate <= step_budget:
            frames = f
            break
    Ls, sgrid = frames * gh * gw, (frames, gh, gw)
    est = _block_flops(cfg, Ls) / rate
    n_timed = max(1, min(int(steps), int(total_budget / max(est, 1e-3))))
    times = [run_block(frames) for _ in range(n_timed)]
    t_step = sum(times) / len(times)
    n_blocks = cfg["num_layers"] + len(range(0, cfg["num_layers"], 2))
    B = 2
    L_full = grid[0] * gh * gw
    scale_up = _block_flops(cfg, L_full) / _block_flops(cfg, Ls)
    s_denoise = t_step * scale_up * n_blocks * B
    tokens_per_s = B * L_full / s_denoise
    cb = dict(value=tokens_per_s, unit="tokens/s", cores=cores, kind=kind,
              sample=f"{what} fp32 on {cores} threads; one step = 1 block of {n_blocks} at dim {D} on 1 sample of the CFG "
                     f"pair, {Ls} tokens = {sgrid} of the {grid[0]}x{gh}x{gw} grid (frames from a 1-frame calibration for "
                     f"<= {step_budget:.0f} s/step, independent of --steps); {len(times)} timed step(s), {t_step:.2f} s/step, "
                     f"{_block_flops(cfg, Ls) / t_step / 1e12:.2f} TFLOP/s; full-grid block = x{scale_up:.2f} by flop count; "
                     f"denoise step = x{n_blocks} blocks x{B} CFG = {s_denoise:.0f} s; tokens/s = {B}*{L_full}/that")
    return cb, t_step, s_denoise, len(times)


def run_reference_arm(args):
    """`bench.py --impl reference`: full-grid blocks on the host cores (the bounded sample of cpu_reference_sample — the same
    one whatever --steps / --warmup say); `steps` in the line = blocks actually timed (<= --steps, capped by the time budget),
    `ms_per_step` the measured time of one (so steps x ms_per_step is the timed wall clock), `value` the metric extrapolated
    from it as cpu_baseline.sample states."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cb, t_step, s_denoise, n_timed = cpu_reference_sample(args.config, steps=max(1, args.steps))
    line = {"impl": "reference", "metric": "video_tokens_per_s", "value": cb["value"], "unit": "tokens/s",
            "n_gpus": args.gpus, "steps": n_timed, "warmup": 0, "requested": {"steps": args.steps, "warmup": args.warmup},
            "ms_per_step": t_step * 1e3,
            "s_per_denoise_step": s_denoise, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": bench_config(args.config, int(os.environ.get("WORLD_SIZE", "1"))),
            "reference_run": "host CPU; ms_per_step = one bounded-sample step (see cpu_baseline.sample), "
                             "s_per_denoise_step / value = extrapolated whole step",
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------------------------
def summarize_timing(timing, peaks):
    """Per kernel class: launches, mean device ms per launch, achieved TFLOP/s or GB/s (work = algorithmic flops/bytes)."""
    out = {}
    for name, recs in timing.items():
        ms = [a.elapsed_time(b) for a, b, _ in recs]
        work = sum(w for _, _, w in recs)
        tot = sum(ms)
        if tot <= 0:
            continue
        ent = {"launches": len(recs), "ms_total": tot, "ms_per_launch": tot / len(recs)}
        if name in ("gemm", "attn_self", "attn_cross", "patch_embed"):
            ent["tflops"] = work / tot / 1e9
            ent["frac_of_bf16_sustained"] = ent["tflops"] / peaks["bf16_tflops_sustained"]
        elif name == "barrier":   # cross-GPU stream barriers (N > 1): time on THIS rank = barrier cost + waiting for the peers
            pass
        else:
            ent["gbs"] = work / tot / 1e6
            ent["frac_of_hbm"] = ent["gbs"] / peaks["hbm_gbs"]
        out[name] = ent
    return out


def run_gpu_arm(args):
    import torch.distributed as dist
    from versecrafter_b200 import _lib, ops
    from versecrafter_b200.transformer import VerseCrafterWanTransformer3DModel

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the B200 path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    lib = _lib.load()
    _lib.check(lib.vc_device_ok(), "vc_device_ok")
    peaks = load_peaks()
    cfg, grid, desc = CONFIGS[args.config]
    B = 2
    L = grid[0] * (grid[1] // 2) * (grid[2] // 2)

    model = VerseCrafterWanTransformer3DModel(**cfg, device="cuda", dtype=torch.bfloat16, init=False)
    model.synthetic_init_(seed=0)
    model.cache_step_invariants = False   # every timed step is a full reference forward (text MLP + 60x cross K/V included)
    model.eval()
    if world > 1:
        model.enable_multi_gpus_inference()
    host = make_inputs(cfg, grid, B)
    dev = dict(x=host["x"].cuda(), geo=host["geo"].cuda(), ctx=[c.cuda() for c in host["ctx"]], t=host["t"].cuda())

    def step_resident():
        return model(x=dev["x"], t=dev["t"], geoada_context=dev["geo"], context=dev["ctx"], seq_len=host["seq_len"],
                     geoada_context_scale=1.0)

    def step_e2e():
        x = host["x"].cuda(non_blocking=True)
        geo = host["geo"].cuda(non_blocking=True)
        ctx = [c.cuda(non_blocking=True) for c in host["ctx"]]
        t = host["t"].cuda(non_blocking=True)
        y = model(x=x, t=t, geoada_context=geo, context=ctx, seq_len=host["seq_len"], geoada_context_scale=1.0)
        # device -> host read of the step's result into a pinned buffer allocated once (a fresh pageable tensor per step
        # would add its allocation and first-touch page faults to every step)
        if host.get("y") is None or host["y"].shape != y.shape:
            host["y"] = torch.empty(y.shape, dtype=y.dtype, pin_memory=True)
        host["y"].copy_(y, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return host["y"]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            out = fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), out

    for _ in range(args.warmup):
        y = step_resident()
    torch.cuda.synchronize()
    assert torch.isfinite(y.float()).all(), "non-finite prediction"

    sp_parity = None
    if world > 1:
        # The sequence-parallel answer must be the single-GPU answer: every rank holds the full weights, so each one
        # computes the P=1 forward of the same inputs locally and compares it with the gathered SP forward at the REAL
        # shape (c3 at P=8: Lp=4095, 5 heads per rank).  Worst rank is reported; a mismatch aborts before timing.
        y_sp = step_resident().float()
        saved = (model.sp_world_size, model.sp_world_rank)
        model.sp_world_size, model.sp_world_rank = 1, 0
        y_one = step_resident().float()
        model.sp_world_size, model.sp_world_rank = saved
        cos = torch.nn.functional.cosine_similarity(y_sp.flatten().double(), y_one.flatten().double(), dim=0)
        rel = (y_sp - y_one).double().norm() / y_one.double().norm()
        stats = torch.tensor([float(cos), -float(rel), 1.0 if torch.equal(y_sp, y_one) else 0.0], device="cuda", dtype=torch.float64)
        dist.all_reduce(stats, op=dist.ReduceOp.MIN)
        sp_parity = {"cosine": float(stats[0]), "rel_l2": -float(stats[1]), "bit_identical": bool(stats[2] == 1.0),
                     "mode": "fused" if model._px is not None else "nccl",
                     "schedule": "interleaved (two chains, split signal/wait barriers)" if model._interleave_active() else "plain",
                     "what": "gathered SP forward vs the rank-local P=1 forward of the same inputs, worst over ranks"}
        del y_sp, y_one
        model.release_workspace()       # the P=1 workspace is not part of the SP step
        torch.cuda.empty_cache()
        assert sp_parity["cosine"] >= 0.9999, f"sequence-parallel forward differs from the single-GPU forward: {sp_parity}"
        for _ in range(2):
            step_resident()             # re-warm after the workspace release

    sampler = ClockSampler(local_rank)   # every rank samples its own GPU; rank 0's goes into `clocks`, all into `per_rank`
    sampler.start()
    # headline: K steps with nothing but the kernels on the stream; the per-kernel CUDA events (two per launch, ~2,000 per
    # step) go into a separate instrumented pass right after, inside the same clock-sampled region
    ops.launch_count = 0
    ops.timing = None
    ms_total, y = timed(step_resident, args.steps)
    launches = ops.launch_count
    ops.timing = {}
    ms_inst, _ = timed(step_resident, args.steps)   # same K steps again, every launch bracketed by CUDA events
    timing, ops.timing = ops.timing, None
    clocks = sampler.stop()
    # host side: wall time the Python / ctypes path needs to ENQUEUE one forward on an idle stream (the launch queue holds
    # the ~1,000 launches of a step, so the call returns before the GPU is done) — the host is the limit only if this
    # approaches the device time of a step
    torch.cuda.synchronize()
    t_h = time.perf_counter()
    step_resident()
    host_ms = (time.perf_counter() - t_h) * 1e3
    torch.cuda.synchronize()
    # end to end through the public forward() with host buffers
    step_e2e()
    ms_e2e, y_host = timed(step_e2e, args.steps)

    ms_step = ms_total / args.steps
    tokens = B * L
    value = tokens / (ms_step / 1e3)
    h2d = host["x"].numel() * 2 + host["geo"].numel() * 2 + sum(c.numel() * 2 for c in host["ctx"]) + host["t"].numel() * 4
    d2h = y_host.numel() * y_host.element_size()
    kern = summarize_timing(timing, peaks)
    per_rank = None
    if world > 1:
        # which GPU the others wait for: device ms per step of every kernel class on every rank (instrumented pass) and each
        # GPU's median SM clock.  Sequence parallelism runs in lock step (two barriers per block), so the step is the
        # slowest rank's compute; a rank's `barrier` time is its own wait for that rank.
        names = sorted(kern)
        mine = torch.tensor([kern[n]["ms_total"] / args.steps for n in names] + [float(clocks["sm_mhz"] or 0.0)],
                            device="cuda", dtype=torch.float64)
        allr = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        tab = torch.stack(allr).cpu()
        per_rank = {n: [round(float(v), 2) for v in tab[:, i]] for i, n in enumerate(names)}
        per_rank["sm_mhz"] = [float(v) for v in tab[:, -1]]
    gemm_f, attn_f = work_per_step(cfg, L, B)
    dom = kern.get("attn_self", {})
    # dominant kernel: self-attention (52 % of the step's FLOPs at c3); algorithmic flops per launch = 4*B*(H/P)*L^2*128
    flops_launch = 4.0 * B * (cfg["num_heads"] / world) * L * L * 128
    achieved = flops_launch / (dom.get("ms_per_launch", float("inf")) * 1e9) if dom else None
    traffic = None
    try:  # per-launch DRAM bytes of the dominant kernel from the committed ncu --set full capture of this config
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            ent = json.load(f).get(f"{args.config}:{world}")
        if ent:
            traffic = ent["dram_bytes_read"] + ent["dram_bytes_write"]
    except Exception:  # noqa: BLE001
        pass
    roofline = {"kernel": "flash_attn_sa_kernel (self-attention)", "bound": "tensor", "achieved": achieved,
                "peak": peaks["bf16_tflops_sustained"], "unit": "TFLOP/s",
                "frac": achieved / peaks["bf16_tflops_sustained"] if achieved else None,
                "peak_source": f"MEASURED_PEAKS.json bf16_tflops_sustained ({peaks['source']}); kernel timed inside a long step",
                "traffic": traffic, "traffic_unit": "bytes/launch (ncu dram__bytes_read.sum + dram__bytes_write.sum)",
                "flops_per_launch": flops_launch, "ms_per_launch": dom.get("ms_per_launch")}
    line = {"metric": "video_tokens_per_s", "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "s_per_denoise_step": ms_step / 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": bench_config(args.config, world),
            "step_pflop": (gemm_f + attn_f) / 1e15,
            "step_frac_of_bf16_sustained": (gemm_f + attn_f) / world / (ms_step * 1e9) / peaks["bf16_tflops_sustained"],
            "roofline": roofline, "kernels": kern,
            "kernels_pass": {"steps": args.steps, "ms_per_step": ms_inst / args.steps,
                             "what": "the same K steps repeated with two CUDA events around every launch; `kernels`, `roofline` "
                                     "come from this pass, `value` / `ms_per_step` from the uninstrumented pass before it"},
            "clocks": clocks,
            "e2e": {"value": tokens / (ms_e2e / args.steps / 1e3), "unit": "tokens/s", "ms_per_step": ms_e2e / args.steps,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": launches, "host_enqueue_ms_per_step": host_ms}
    if sp_parity is not None:
        line["sp_parity"] = sp_parity
        line["per_rank_ms_per_step"] = per_rank
    if world == 1 and not args.no_gpu_reference:
        # the bar on the same box (BASELINE.md §4 B-GPU-ref): one block of the reference's own module (autocast bf16 +
        # flash-attn) next to the same block on this repo's kernels, at this config's shape — after every timed region
        try:
            model.release_workspace()
            torch.cuda.empty_cache()
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            from ref_gpu_block import measure_block
            line["gpu_reference"] = measure_block(model, cfg, (grid[0], grid[1] // 2, grid[2] // 2), iters=3, B=B)
        except Exception as e:  # noqa: BLE001
            line["gpu_reference"] = {"unavailable": repr(e)[:300]}
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            try:
                line["cpu_baseline"] = cpu_reference_sample(args.config, steps=2)[0]
            except Exception as e:  # noqa: BLE001
                line["cpu_baseline"] = {"error": repr(e)[:200]}
        print(json.dumps(line), flush=True)
    if world > 1:
        if model._peer_exchange() is not None:
            model._peer_exchange().check_status(wait=True)   # no cross-GPU barrier of the run timed out
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default=os.environ.get("VC_BENCH_CONFIG", "c3"), choices=sorted(CONFIGS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-reference", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
