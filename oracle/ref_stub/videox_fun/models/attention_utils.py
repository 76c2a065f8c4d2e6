"""The one arithmetic substitution of the oracle: VideoX-Fun's `attention()` (un-vendored) is bound to
  * flash_attn.flash_attn_func for bf16/fp16 CUDA tensors when flash_attn is importable — the kernel VideoX-Fun's own
    attention() dispatches to on a GPU, which makes the imported reference the "PyTorch + flash-attn path" of north_star;
  * otherwise softmax(QK^T/sqrt(d))V with no mask, computed by torch SDPA in fp32 (CPU fixtures, fp32 runs).
No mask either way: every parity config has k_lens == L and cross-attention passes k_lens=None."""
import os

import torch
import torch.nn.functional as F

USE_FLASH = os.environ.get("VC_REF_ATTENTION", "auto")  # auto | flash | sdpa
LAST_IMPL = None  # what the most recent call used (tests assert on it)


def _flash():
    if USE_FLASH == "sdpa":
        return None
    try:
        from flash_attn import flash_attn_func
        return flash_attn_func
    except Exception:  # noqa: BLE001
        if USE_FLASH == "flash":
            raise
        return None


def attention(q, k, v, q_lens=None, k_lens=None, dropout_p=0.0, softmax_scale=None, q_scale=None,
              causal=False, window_size=(-1, -1), deterministic=False, dtype=torch.bfloat16, **kw):
    global LAST_IMPL
    out_dtype = q.dtype
    if q.is_cuda and q.dtype in (torch.bfloat16, torch.float16):
        fa = _flash()
        if fa is not None:
            LAST_IMPL = "flash_attn"
            return fa(q.contiguous(), k.contiguous(), v.contiguous(), softmax_scale=softmax_scale)
    LAST_IMPL = "sdpa_fp32"
    q, k, v = (u.transpose(1, 2).float() for u in (q, k, v))
    o = F.scaled_dot_product_attention(q, k, v, scale=softmax_scale)
    return o.transpose(1, 2).contiguous().to(out_dtype)
