"""CPU checks of the drop-in boundary: the C-ABI library builds/loads without a GPU and exports every symbol that
include/vc_b200.h declares; the host-side module mirrors the reference's parameter names; no CPU fallback exists."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from versecrafter_b200 import build, _lib
    build.build()
    return _lib.load()


def test_library_exports_every_declared_symbol(lib):
    hdr = open(os.path.join(ROOT, "include", "vc_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(vc_[a-z0-9_]+)\s*\(", hdr))
    assert {"vc_gemm_bf16", "vc_flash_attn_fwd", "vc_ln_modulate", "vc_rmsnorm_rope_qk", "vc_patchify"} <= declared
    raw = ctypes.CDLL(os.path.join(ROOT, "versecrafter_b200", "libvc_b200.so"))
    for name in declared:
        assert hasattr(raw, name), f"{name} declared in include/vc_b200.h but not exported"
    from versecrafter_b200._lib import SIGNATURES
    assert set(SIGNATURES) == declared


def test_ctypes_table_matches_the_header_prototypes():
    """Every prototype of include/vc_b200.h against versecrafter_b200/_lib.SIGNATURES: same number of arguments, and each
    argument of the same kind (pointer / 64-bit integer / 32-bit integer / float) — a drift here would not fail to load, it
    would pass garbage to a kernel launcher."""
    from versecrafter_b200._lib import SIGNATURES
    hdr = open(os.path.join(ROOT, "include", "vc_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    hdr = "\n".join(l for l in hdr.splitlines() if not l.lstrip().startswith("#"))
    protos = re.findall(r"([A-Za-z_][A-Za-z0-9_ \*]*?)\b(vc_[a-z0-9_]+)\s*\(([^)]*)\)\s*;", hdr, flags=re.S)
    assert len(protos) == len(SIGNATURES)

    def kind_c(decl):
        decl = " ".join(decl.split())
        if decl == "void" or not decl:
            return None
        if "*" in decl:
            return "ptr"
        base = decl.rsplit(" ", 1)[0] if " " in decl else decl
        if "long long" in base:
            return "i64"
        if "float" in base:
            return "f32"
        if "double" in base:
            return "f64"
        if "int" in base or "unsigned" in base:
            return "i32"
        raise AssertionError(f"unclassified parameter {decl!r}")

    def kind_ct(t):
        if t in (ctypes.c_void_p, ctypes.c_char_p) or hasattr(t, "_type_") and not isinstance(t._type_, str):
            return "ptr"
        return {ctypes.c_longlong: "i64", ctypes.c_int: "i32", ctypes.c_uint: "i32", ctypes.c_float: "f32",
                ctypes.c_double: "f64"}[t]

    for ret, name, args in protos:
        want = [k for k in (kind_c(a) for a in args.split(",")) if k]
        res, argtypes = SIGNATURES[name]
        got = [kind_ct(t) for t in argtypes]
        assert got == want, f"{name}: header {want} vs ctypes {got}"
        assert (kind_ct(res) == "ptr") == ("*" in ret), name


def test_version_and_arch(lib):
    assert lib.vc_version() >= 100
    assert lib.vc_build_arch() == 100  # sm_100a only


def test_argument_errors_do_not_need_a_gpu(lib):
    rc = lib.vc_gemm_bf16(None, 8, None, 8, None, 8, 1, 8, 8, 0, None, None, 0, 0, None, 0, 1.0, None)
    assert rc < 0 and b"null" in lib.vc_last_error()
    # A overlapping C is rejected (the persistent kernel re-reads A rows after C tiles of those rows are stored) ...
    P = ctypes.c_void_p
    rc = lib.vc_gemm_bf16(P(4096), 512, P(1 << 30), 512, P(4096), 512, 300, 512, 512, 0, None, None, 0, 0, None, 0, 1.0, None)
    assert rc < 0 and b"overlap" in lib.vc_last_error()
    rc = lib.vc_gemm_bf16(P(4096), 512, P(1 << 30), 512, P(4096 + 600 * 2), 1024, 300, 512, 512, 0, None, None, 0, 0, None, 0, 1.0, None)
    assert rc < 0 and b"overlap" in lib.vc_last_error()
    # ... but disjoint column blocks of one row-strided buffer (A = buf[:, :512], C = buf[:, 512:1024]) are not
    # (fake addresses: only where no device exists, so the call stops at the tensor-map encoding instead of launching)
    if not torch.cuda.is_available():
        rc = lib.vc_gemm_bf16(P(4096), 1536, P(1 << 30), 512, P(4096 + 1024), 1536, 300, 512, 512, 0, None, None, 0, 0, None, 0,
                              1.0, None)
        assert rc != 0 and b"overlap" not in lib.vc_last_error()
    rc = lib.vc_ln_modulate(ctypes.c_void_p(16), 12, ctypes.c_void_p(16), 12, ctypes.c_void_p(16), ctypes.c_void_p(16), 0, 4, 4,
                            12, 1e-6, 0, None)
    assert rc < 0 and b"multiple of 8" in lib.vc_last_error()
    # split barrier: rank outside the group, null flag table
    flags = (ctypes.c_void_p * 8)(*[4096] * 8)
    assert lib.vc_peer_signal(flags, 9, 8, 1, None) < 0 and b"bad rank" in lib.vc_last_error()
    assert lib.vc_peer_wait(None, 0, 2, 1, 0, ctypes.c_void_p(64), None) < 0 and b"null" in lib.vc_last_error()
    assert lib.vc_peer_wait(flags, 0, 9, 1, 0, ctypes.c_void_p(64), None) < 0 and b"bad rank" in lib.vc_last_error()


def test_sass_is_blackwell_native():
    """tcgen05.mma / tcgen05.ld / TMA must be in the shipped SASS (UTCHMMA / LDTM / UTMALDG), no legacy HMMA."""
    import shutil
    import subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    sass = subprocess.run([cuobjdump, "-sass", os.path.join(ROOT, "versecrafter_b200", "libvc_b200.so")],
                          capture_output=True, text=True).stdout
    assert "UTCHMMA" in sass and "LDTM" in sass and "UTMALDG" in sass
    assert "HMMA.16816" not in sass
    assert "sm_100a" in sass or "SM100" in sass.upper()


def test_no_cpu_fallback():
    from versecrafter_b200 import _lib, ops
    from versecrafter_b200.transformer import VerseCrafterWanTransformer3DModel
    a = torch.zeros(8, 8, dtype=torch.bfloat16)
    with pytest.raises(_lib.VCError):
        ops.gemm(a, a, None)
    m = VerseCrafterWanTransformer3DModel(dim=128, ffn_dim=256, num_heads=1, num_layers=2, text_dim=32, geoada_in_dim=128)
    m = m.to(torch.bfloat16)
    with pytest.raises(_lib.VCError):
        m(torch.zeros(1, 16, 1, 4, 4), torch.zeros(1), torch.zeros(1, 128, 1, 4, 4), [torch.zeros(3, 32)], 4)


def test_state_dict_keys_match_oracle_and_reference():
    from oracle import ref_import, restate
    from versecrafter_b200.transformer import VerseCrafterWanTransformer3DModel
    cfg = dict(dim=128, ffn_dim=256, num_heads=1, num_layers=4, text_dim=32, geoada_in_dim=128)
    m = VerseCrafterWanTransformer3DModel(**cfg)
    shapes = {k: tuple(v.shape) for k, v in m.state_dict().items()}
    assert shapes == {k: tuple(v) for k, v in restate.model_param_shapes(cfg).items()}
    if ref_import.reference_available():
        _, vc = ref_import.load_reference_modules()
        ref = vc.VerseCrafterWanTransformer3DModel(**cfg)
        assert shapes == {k: tuple(v.shape) for k, v in ref.state_dict().items()}
        assert [b.block_id for b in m.blocks] == [b.block_id for b in ref.blocks]
        assert [b.block_id for b in m.geoada_blocks] == [b.block_id for b in ref.geoada_blocks]
        assert torch.equal(m.freqs, ref.freqs)
        # fresh reference-style init reproduces the reference's trivially-valued parameters
        for k, v in ref.state_dict().items():
            if (v == 0).all() or (v == 1).all():
                assert torch.equal(m.state_dict()[k].float(), v.float()), k


def test_packed_weights_follow_load_state_dict_and_to():
    from oracle import restate
    from versecrafter_b200.transformer import VerseCrafterWanTransformer3DModel
    cfg = dict(dim=128, ffn_dim=256, num_heads=1, num_layers=2, text_dim=32, geoada_in_dim=128)
    m = VerseCrafterWanTransformer3DModel(**cfg).to(torch.bfloat16)
    blk = m.blocks[1]
    pk = blk.ensure_packed()
    sd = restate.synth_state_dict(cfg, seed=3)
    m.load_state_dict(sd)
    assert torch.equal(pk["wqkv"], torch.cat([sd[f"blocks.1.self_attn.{n}.weight"] for n in "qkv"]))
    assert torch.equal(pk["bkv"], torch.cat([sd[f"blocks.1.cross_attn.{n}.bias"] for n in "kv"]))
    m = m.to(torch.float32).to(torch.bfloat16)  # .to() re-allocates every parameter: views must be rebuilt
    pk2 = m.blocks[1].ensure_packed()
    assert pk2 is not pk and torch.equal(pk2["wqkv"], pk["wqkv"])
    assert m.blocks[1].self_attn.k.weight.data_ptr() == pk2["wqkv"][128:].data_ptr()
    assert set(m.state_dict()) == set(sd)


def test_rope_table_matches_oracle_angles():
    from oracle import restate
    from versecrafter_b200 import ops
    grid = (3, 5, 7)
    table, (nf, nh, nw) = ops.rope_table(128, grid, "cpu")
    ang = restate.rope_angles(128, grid)
    assert (nf, nh, nw) == (22, 21, 21)
    tf = table[:grid[0] * nf].view(grid[0], nf, 2)
    tw = table[grid[0] * nf + grid[1] * nh:].view(grid[2], nw, 2)
    n = (2 * grid[1] + 3) * grid[2] + 4  # token (f=2, h=3, w=4)
    assert torch.allclose(tf[2, :, 0].double(), ang[n, :nf].cos(), atol=1e-7)
    assert torch.allclose(tw[4, :, 1].double(), ang[n, nf + nh:].sin(), atol=1e-7)


def test_riflex_and_teacache_surface():
    from versecrafter_b200.transformer import VerseCrafterWanTransformer3DModel
    m = VerseCrafterWanTransformer3DModel(dim=128, ffn_dim=256, num_heads=1, num_layers=2, text_dim=32, geoada_in_dim=128)
    f0 = m.freqs.clone()
    m.enable_riflex(k=6, L_test=30)
    assert not torch.equal(m.freqs, f0) and torch.equal(m.freqs[:, 22:], f0[:, 22:])
    m.disable_riflex()
    assert torch.equal(m.freqs, f0)
    m.enable_teacache([1.0, 0.0], 10, 0.1, num_skip_start_steps=2, offload=False)
    assert m.teacache.cnt == 0 and m.teacache.rel_l1_thresh == 0.1
    m.enable_cfg_skip(0.25, 40)
    assert m.cfg_skip_ratio == 0.25 and m.num_inference_steps == 40
    m.disable_cfg_skip(); m.disable_teacache()
    assert m.teacache is None and m.cfg_skip_ratio is None
