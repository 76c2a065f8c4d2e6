"""Ulysses sequence parallelism around self-attention: every op of the DiT block is token-local
except self-attention, so each rank owns a contiguous chunk of Lp = L/P tokens of every sample (the reference's
torch.chunk(dim=1), wan_transformer3d_versecrafter.py:269-270, :366-367) and swaps "all heads of my tokens" for
"my heads of all tokens" with one all-to-all before attention and the inverse after it — what the reference gets
from xfuser's SeqAllToAll4D inside usp_attn_forward (not in the reference tree; ring attention is dropped by design).

The layout transforms are plain tensor permutations so they run on any backend: NCCL over NVLink on the B200 box,
gloo on CPU in tests/test_sp_cpu.py.
"""
import torch
import torch.distributed as dist


def scatter_heads(x, P, group=None):
    """x [B, Lp, H, d] (all heads, my tokens) -> [B, P*Lp, H/P, d] (my heads, all tokens; token order = rank order)."""
    B, Lp, H, d = x.shape
    Hp = H // P
    # send[p] = my tokens, heads of rank p
    send = x.reshape(B, Lp, P, Hp, d).permute(2, 0, 1, 3, 4).contiguous()      # [P, B, Lp, Hp, d]
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv, send, group=group)                            # recv[p] = tokens of rank p, my heads
    return recv.permute(1, 0, 2, 3, 4).reshape(B, P * Lp, Hp, d)              # one more copy: [B, L, Hp, d]


def gather_heads(o, P, group=None):
    """o [B, P*Lp, H/P, d] (my heads, all tokens) -> [B, Lp, H, d] (all heads, my tokens)."""
    B, L, Hp, d = o.shape
    Lp = L // P
    send = o.reshape(B, P, Lp, Hp, d).permute(1, 0, 2, 3, 4).contiguous()      # [P, B, Lp, Hp, d]: tokens of rank p
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv, send, group=group)                            # recv[p] = my tokens, heads of rank p
    return recv.permute(1, 2, 0, 3, 4).reshape(B, Lp, P * Hp, d)


def all_gather_tokens(x, dim, group=None):
    """Concatenate the rank chunks along `dim` (get_sp_group().all_gather, wan_transformer3d_versecrafter.py:432-433)."""
    P = dist.get_world_size(group)
    parts = [torch.empty_like(x) for _ in range(P)]
    dist.all_gather(parts, x.contiguous(), group=group)
    return torch.cat(parts, dim=dim)


def ulysses_attention(qkv, out, B, Lp, H, d, P, group, buf, variant=0):
    """qkv [B*Lp, 3*H*d] (q,k already normalised + rotated with GLOBAL positions) -> out [B*Lp, H*d]."""
    from . import ops
    D = H * d
    q, k, v = (scatter_heads(qkv[:, i * D:(i + 1) * D].reshape(B, Lp, H, d), P, group) for i in range(3))
    o = ops.flash_attn(q, k, v, variant=variant)
    out.view(B, Lp, H, d).copy_(gather_heads(o, P, group))
    return out
