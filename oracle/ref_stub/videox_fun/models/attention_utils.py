"""The one arithmetic substitution of the oracle: VideoX-Fun's `attention()` is bound to
softmax(QK^T/sqrt(d))V with no mask, computed by torch SDPA in the input dtype's fp32 upcast."""
import torch
import torch.nn.functional as F


def attention(q, k, v, q_lens=None, k_lens=None, dropout_p=0.0, softmax_scale=None, q_scale=None,
              causal=False, window_size=(-1, -1), deterministic=False, dtype=torch.bfloat16, **kw):
    out_dtype = q.dtype
    q, k, v = (u.transpose(1, 2).float() for u in (q, k, v))
    o = F.scaled_dot_product_attention(q, k, v, scale=softmax_scale)
    return o.transpose(1, 2).contiguous().to(out_dtype)
