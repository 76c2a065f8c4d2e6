"""TEST INFRASTRUCTURE — loads the reference's own two model files, unmodified.

Source: /root/reference where it exists (the build container), else the byte-for-byte copy that oracle/build_ref.py
placed under the git-ignored oracle/_ref/ (it travels with the gpurun snapshot; digests pinned in
oracle/ref_manifest.json), so the reference itself can run on the B200 under torch.autocast(bf16) with attention()
bound to flash_attn — the "reference PyTorch + flash-attn path" north_star names.  Used by tests/golden/gen_golden.py
to produce the committed fixtures, by the CPU tests to pin oracle/restate.py's fp32 mode, by tests/test_reference_gpu.py
to pin its bf16 mode and the drop-in against the reference, and by bench.py's CPU arm.

The import goes through oracle/ref_stub (fake `diffusers`, fake `videox_fun`) because neither package is
installed and VideoX-Fun is an un-vendored submodule with no recoverable pin.  The only
arithmetic substitution is `attention()` -> torch SDPA (oracle/ref_stub/videox_fun/models/attention_utils.py).
"""
import importlib.util
import os
import sys

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_STUB = os.path.join(_HERE, "ref_stub")
_PROBE = "versecrafter/models/wan_transformer3d_versecrafter.py"


def reference_root():
    """/root/reference (or $VC_REFERENCE_ROOT) when present, else oracle/_ref if it verifies against the pinned
    digests, else None."""
    live = os.environ.get("VC_REFERENCE_ROOT", "/root/reference")
    if os.path.isfile(os.path.join(live, _PROBE)):
        return live
    local = os.path.join(_HERE, "_ref")
    if os.path.isfile(os.path.join(local, _PROBE)):
        from . import build_ref
        if build_ref.verify(local):
            return local
    return None


REF_ROOT = reference_root() or "/root/reference"


def reference_available():
    return os.path.isfile(os.path.join(REF_ROOT, _PROBE))


def _load(name, path):
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def load_reference_modules():
    """Returns (wan_transformer3d, wan_transformer3d_versecrafter) modules of the reference."""
    if not reference_available():
        raise FileNotFoundError(f"reference not found under {REF_ROOT}")
    if _STUB not in sys.path:
        sys.path.insert(0, _STUB)
    # Parent packages as empty namespaces so the relative import `.wan_transformer3d` resolves without
    # executing versecrafter/models/__init__.py (which needs the real diffusers/transformers classes).
    import types
    for pkg in ("versecrafter", "versecrafter.models"):
        if pkg not in sys.modules:
            m = types.ModuleType(pkg)
            m.__path__ = []
            sys.modules[pkg] = m
    base = _load("versecrafter.models.wan_transformer3d",
                 os.path.join(REF_ROOT, "versecrafter/models/wan_transformer3d.py"))
    vc = _load("versecrafter.models.wan_transformer3d_versecrafter",
               os.path.join(REF_ROOT, "versecrafter/models/wan_transformer3d_versecrafter.py"))
    return base, vc


def perturb_trivial_params_(state_or_model, seed=1234):
    """a freshly built model outputs exactly 0 (zero-init head / after_proj /
    before_proj) and has unit gains / zero biases that would hide dropped terms.  Re-initialise every
    trivially valued parameter so parity is not vacuous.  Works on an nn.Module or a state dict."""
    sd = state_or_model if isinstance(state_or_model, dict) else dict(state_or_model.named_parameters())
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name in sorted(sd):
            p = sd[name]
            is_gain = name.endswith(("norm_q.weight", "norm_k.weight", "norm3.weight"))
            trivially_zero = bool((p == 0).all())
            trivially_one = bool((p == 1).all())
            if is_gain and trivially_one:
                p.copy_(1.0 + 0.1 * torch.randn(p.shape, generator=g))
            elif trivially_zero:
                p.copy_(0.02 * torch.randn(p.shape, generator=g))
    return state_or_model


def build_reference_model(cfg, seed=0):
    """cfg: dict of ctor kwargs for VerseCrafterWanTransformer3DModel
    (/root/reference/versecrafter/models/wan_transformer3d_versecrafter.py:152-170)."""
    _, vc = load_reference_modules()
    torch.manual_seed(seed)
    model = vc.VerseCrafterWanTransformer3DModel(**cfg)
    perturb_trivial_params_(model, seed=seed + 1234)
    model.eval().requires_grad_(False)
    return model
