"""ctypes binding of libvc_b200.so (include/vc_b200.h).  There is no fallback: if the library is missing or a call
fails, an exception is raised."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvc_b200.so")

_c_ll = ctypes.c_longlong
_vp = ctypes.c_void_p
_i = ctypes.c_int
_f = ctypes.c_float
_llp = ctypes.POINTER(ctypes.c_longlong)
_vpp = ctypes.POINTER(ctypes.c_void_p)
_fp = ctypes.POINTER(ctypes.c_float)

SIGNATURES = {
    "vc_version": (_i, []),
    "vc_build_arch": (_i, []),
    "vc_device_ok": (_i, []),
    "vc_last_error": (ctypes.c_char_p, []),
    "vc_ln_modulate": (_i, [_vp, _c_ll, _vp, _c_ll, _vp, _vp, _c_ll, _c_ll, _c_ll, _i, _f, _i, _vp]),
    "vc_rmsnorm_rope_qk": (_i, [_vp, _c_ll, _c_ll, _c_ll, _c_ll, _i, _i, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _f, _i, _vp]),
    "vc_gemm_bf16": (_i, [_vp, _c_ll, _vp, _c_ll, _vp, _c_ll, _c_ll, _c_ll, _c_ll, _i, _vp, _vp, _c_ll, _c_ll, _vp, _c_ll, _f, _vp]),
    "vc_flash_attn_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _llp, _llp, _llp, _llp, _f, _i, _vp]),
    "vc_rmsnorm_rope_qkv_scatter": (_i, [_vp, _c_ll, _c_ll, _c_ll, _c_ll, _i, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _f,
                                         _vpp, _i, _i, _c_ll, _i, _vp]),
    "vc_peer_barrier": (_i, [_vpp, _i, _i, ctypes.c_uint, _i, _vp, _vp]),
    "vc_peer_signal": (_i, [_vpp, _i, _i, ctypes.c_uint, _vp]),
    "vc_peer_wait": (_i, [_vpp, _i, _i, ctypes.c_uint, _i, _vp, _vp]),
    "vc_gemm_bf16_scatter": (_i, [_vp, _c_ll, _vp, _c_ll, _c_ll, _c_ll, _c_ll, _vp, _c_ll, _c_ll, _vpp, _i, _i, _i, _c_ll, _i,
                                  _vp]),
    "vc_flash_attn_fwd_scatter": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _llp, _llp, _llp, _vpp, _i, _i, _llp, _f, _i,
                                       _vp]),
    "vc_patch_embed_supported": (_i, [_i, _i, _i, _i, _i]),
    "vc_patch_embed": (_i, [_vp, _i, _i, _i, _i, _vp, _c_ll, _vp, _i, _vp, _c_ll, _c_ll, _c_ll, _vp]),
    "vc_control_latents": (_i, [_vpp, _i, _i, _vp, _i, _i, _i, _i, _i, ctypes.POINTER(ctypes.c_int), _vp, _vp]),
    "vc_cfg_euler_step": (_i, [_vp, _vp, _vp, _vp, _vp, _c_ll, _f, _f, _vp]),
    "vc_cfg_unipc_step": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _c_ll, _f, _fp, _vp]),
    "vc_composite_by_depth": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _c_ll, _vp]),
    "vc_depth_to_gray": (_i, [_vp, _vp, _c_ll, _f, _f, _i, _vp]),
    "vc_gaussian_projection": (_i, [_i, ctypes.POINTER(ctypes.c_float), _i, _i, _f, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "vc_render_points": (_i, [_vp, _vp, _c_ll, _fp, _fp, _i, _i, _f, _fp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "vc_render_mesh_depth": (_i, [_vp, _i, _vp, _i, _fp, _fp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "vc_time_embed": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "vc_time_embed_static": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "vc_modulation_table": (_i, [_vp, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "vc_patchify": (_i, [_vp, _vp, _i, _i, _i, _i, _c_ll, _vp]),
    "vc_unpatchify": (_i, [_vp, _c_ll, _vp, _i, _i, _i, _i, _vp]),
}

_lib = None


class VCError(RuntimeError):
    pass


def load():
    """Loads the library (building it is the job of `python -m versecrafter_b200.build` / __graft_entry__.build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise VCError(f"{LIB_PATH} not found: build it with `python -m versecrafter_b200.build` "
                      f"(there is no CPU or PyTorch fallback for the sm_100a kernels)")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().vc_last_error()
        raise VCError(f"{what} failed (rc={rc}): {msg.decode() if msg else ''}")
