"""The one arithmetic substitution of the oracle: VideoX-Fun's `attention()` (un-vendored) is bound to
  * flash_attn.flash_attn_func for bf16/fp16 CUDA tensors when flash_attn is importable — the kernel VideoX-Fun's own
    attention() dispatches to on a GPU, which makes the imported reference the "PyTorch + flash-attn path" of north_star;
  * otherwise softmax(QK^T/sqrt(d))V with no mask, computed by torch SDPA in fp32 (CPU fixtures, fp32 runs).
k_lens (self-attention passes seq_lens, wan_transformer3d.py:394-399) is honoured the way VideoX-Fun's varlen flash path does:
sample b attends to its first k_lens[b] keys only, every query row (padded ones included) is still computed.  With
k_lens == Lk or None (cross-attention, :425-430) there is no mask."""
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
    Lk = k.shape[1]
    lens = None
    if k_lens is not None:
        lens = [min(int(n), Lk) for n in torch.as_tensor(k_lens).reshape(-1).tolist()]
        if all(n == Lk for n in lens):
            lens = None
    if q.is_cuda and q.dtype in (torch.bfloat16, torch.float16):
        fa = _flash()
        if fa is not None:
            LAST_IMPL = "flash_attn"
            if lens is None:
                return fa(q.contiguous(), k.contiguous(), v.contiguous(), softmax_scale=softmax_scale)
            return torch.cat([fa(q[b:b + 1].contiguous(), k[b:b + 1, :n].contiguous(), v[b:b + 1, :n].contiguous(),
                                 softmax_scale=softmax_scale) for b, n in enumerate(lens)])
    LAST_IMPL = "sdpa_fp32"
    q, k, v = (u.transpose(1, 2).float() for u in (q, k, v))
    mask = None
    if lens is not None:   # [B, 1, 1, Lk]: True = key takes part
        mask = (torch.arange(Lk, device=q.device)[None, :] < torch.tensor(lens, device=q.device)[:, None])[:, None, None, :]
    o = F.scaled_dot_product_attention(q, k, v, attn_mask=mask, scale=softmax_scale)
    return o.transpose(1, 2).contiguous().to(out_dtype)
