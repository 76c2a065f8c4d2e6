"""TEST INFRASTRUCTURE — runs the reference's own VerseCrafterWanTransformer3DModel (imported unmodified through
oracle/ref_import.py) the way the pipeline runs it and captures what every block saw and returned, with forward
hooks.  On CUDA: bf16 weights under torch.autocast(bf16) (pipeline_wan_versecrafter.py:893) with attention() bound to
flash_attn; on CPU: plain fp32.  Used by tests/test_reference_gpu.py and tests/test_oracle_golden.py."""
import torch

from . import restate


def run_reference(ref_mods, cfg, grid, seed, scale=1.0, t_value=500.0, device="cuda", dtype=torch.bfloat16, seq_pad=0):
    """-> dict with the reference's bf16 output and, per block, the tensors its forward saw / returned.  seq_pad > 0 asks
    for seq_len = L + seq_pad: the reference zero-pads the token rows and masks them as keys (k_lens = L < seq_len)."""
    base, vc = ref_mods
    sd = restate.synth_state_dict(cfg, seed=seed, device=device, dtype=dtype)
    ref = vc.VerseCrafterWanTransformer3DModel(**cfg)
    ref.load_state_dict(sd, strict=True)
    ref = ref.to(device=device, dtype=dtype).eval().requires_grad_(False)
    inp = restate.synth_inputs(cfg, grid, seed=seed, device=device, t_value=t_value)
    inp["seq_len"] += int(seq_pad)
    cap = {}

    def backbone_hook(i):
        def pre(mod, args, kwargs):
            cap[f"blk_in.{i}"] = args[0].detach().clone()
            cap[f"blk_kwargs.{i}"] = {k: v for k, v in kwargs.items() if k not in ("hints", "context_scale")}
        def post(mod, args, kwargs, out):
            cap[f"blk_out.{i}"] = out.detach().clone()
        return pre, post

    def geo_hook(j):
        def pre(mod, args, kwargs):
            cap[f"geo_stack_in.{j}"] = args[0].detach().clone()
            cap[f"geo_kwargs.{j}"] = dict(kwargs)
        def post(mod, args, kwargs, out):
            cap[f"geo_out.{j}"] = out[-1].detach().clone()
            cap[f"hint.{j}"] = out[-2].detach().clone()
        return pre, post

    handles = []
    for i, blk in enumerate(ref.blocks):
        pre, post = backbone_hook(i)
        handles += [blk.register_forward_pre_hook(pre, with_kwargs=True), blk.register_forward_hook(post, with_kwargs=True)]
    for j, blk in enumerate(ref.geoada_blocks):
        pre, post = geo_hook(j)
        handles += [blk.register_forward_pre_hook(pre, with_kwargs=True), blk.register_forward_hook(post, with_kwargs=True)]
    import contextlib
    amp_ctx = torch.autocast("cuda", dtype=dtype) if torch.device(device).type == "cuda" else contextlib.nullcontext()
    with torch.no_grad(), amp_ctx:
        y = ref(x=inp["x"].to(dtype), t=inp["t"], geoada_context=inp["geoada_context"].to(dtype),
                context=[c.to(dtype) for c in inp["context"]], seq_len=inp["seq_len"], geoada_context_scale=scale)
        # block cores, recomputed by the reference's own WanAttentionBlock.forward on the captured inputs
        for i, blk in enumerate(ref.blocks):
            cap[f"blk_core.{i}"] = base.WanAttentionBlock.forward(blk, cap[f"blk_in.{i}"], **cap[f"blk_kwargs.{i}"])
        for j, blk in enumerate(ref.geoada_blocks):
            kw = cap[f"geo_kwargs.{j}"]
            c = cap[f"geo_stack_in.{j}"]
            cap[f"geo_in.{j}"] = (blk.before_proj(c) + kw["x"]) if j == 0 else c[-1]
    for h in handles:
        h.remove()
    from videox_fun.models import attention_utils
    cap["attention_impl"] = attention_utils.LAST_IMPL
    kw0 = cap["blk_kwargs.0"]
    cap["e0"], cap["ctx"] = kw0["e"], kw0["context"]
    return dict(y=y, cap=cap, sd=sd, inp=inp)
