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


def ulysses_attention(qkv, out, B, Lp, H, d, P, group, buf, variant=0, k_true=None):
    """qkv [B*Lp, 3*H*d] (q,k already normalised + rotated with GLOBAL positions) -> out [B*Lp, H*d].
    k_true: number of real tokens (zero-padded tail rows past it are not attended, as on one GPU)."""
    from . import ops
    D = H * d
    q, k, v = (scatter_heads(qkv[:, i * D:(i + 1) * D].reshape(B, Lp, H, d), P, group) for i in range(3))
    if k_true is not None and k_true < P * Lp:
        k, v = k[:, :k_true], v[:, :k_true]
    o = ops.flash_attn(q, k, v, variant=variant)
    out.view(B, Lp, H, d).copy_(gather_heads(o, P, group))
    return out


# ------------------------------------------------------------------------------------------------------------------
# Fused path: the all-to-alls live inside the producer kernels (stores to peer-mapped memory over NVLink)
# ------------------------------------------------------------------------------------------------------------------
class PeerExchange:
    """Symmetric receive buffers of one sequence-parallel group: every rank allocates the same buffers, the group
    exchanges their device addresses once (torch symmetric memory = CUDA VMM handles mapped into every process), and
    kernels then store straight into the peers' buffers.  `barrier()` is a stream-ordered cross-GPU barrier (signal
    pads in the same symmetric allocation): work enqueued after it on any rank starts only when every rank's stream
    has reached it, i.e. after every peer's earlier kernels (and their remote stores) have completed."""

    def __init__(self, group, device):
        import torch.distributed._symmetric_memory as symm
        self.symm = symm
        self.group = group if group is not None else dist.group.WORLD
        self.P = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.device = device
        self.bufs = {}

    def buffer(self, name, numel):
        """bf16 symmetric buffer of at least `numel` elements -> (tensor, [device address of every rank's copy])."""
        ent = self.bufs.get(name)
        if ent is None or ent[0].numel() < numel:
            t = self.symm.empty(int(numel), dtype=torch.bfloat16, device=self.device)
            h = self.symm.rendezvous(t, self.group.group_name)
            ent = (t, [int(p) for p in h.buffer_ptrs], h)
            self.bufs[name] = ent
        return ent[0], ent[1]

    def barrier(self):
        next(iter(self.bufs.values()))[2].barrier(channel=0)


def recv_layout_index(which, b, t, h, n_batch, l_total, heads_per_peer, head_dim):
    """(peer, element offset) where vc_rmsnorm_rope_qkv_scatter stores head h of global token t of sample b
    (which = 0/1/2 for q/k/v) — the host-side statement of the kernel's destination formula (include/vc_b200.h)."""
    peer, hl = divmod(h, heads_per_peer)
    return peer, ((((which * n_batch + b) * l_total + t) * heads_per_peer) + hl) * head_dim


def attn_layout_index(b, t, h, rows_per_peer, n_heads, head_dim):
    """(peer, element offset) where vc_flash_attn_fwd_scatter stores output row t (global token), global head h."""
    peer, tl = divmod(t, rows_per_peer)
    return peer, ((b * rows_per_peer + tl) * n_heads + h) * head_dim


def ulysses_attention_fused(px, qkv, B, Lp, H, d, gain_q, gain_k, rope, grid, eps, variant, k_true=None):
    """qkv [B*Lp, 3*H*d] straight from the fused QKV GEMM (NOT yet normalised) -> attention output [B*Lp, H*d] (a view
    of this rank's symmetric buffer).  Two kernels, two barriers, no NCCL call and no layout copy:
      1. RMSNorm+RoPE(q,k) | v  -> stored head-scattered into every peer's [3, B, L, H/P, d] buffer (all-to-all #1-3)
      2. flash attention over all L tokens for the local H/P heads -> each output row stored into the owner rank's
         [B, Lp, H, d] buffer (all-to-all #4)."""
    from . import ops
    P, rank = px.P, px.rank
    Hp, L, D = H // P, P * Lp, H * d
    n_recv, n_attn = 3 * B * L * Hp * d, B * Lp * D
    recv_t, recv_ptrs = px.buffer("qkv_recv", n_recv)
    attn_t, attn_ptrs = px.buffer("attn_recv", n_attn)
    ops.rmsnorm_rope_qkv_scatter(qkv, D, gain_q, gain_k, Lp, recv_ptrs, Hp, L, rope, grid, pos_offset=rank * Lp, eps=eps)
    px.barrier()  # every peer's q/k/v slices have landed here; peers are done reading attn_recv of the previous block
    r = recv_t[:n_recv].view(3, B, L, Hp, d)
    head_off = rank * Hp * d * 2  # bytes: this rank's heads inside a peer's [.., H, d] row
    Lk = L if k_true is None else min(k_true, L)  # padded tail tokens are not attended (k_lens, as on one GPU)
    ops.flash_attn_scatter(r[0], r[1][:, :Lk], r[2][:, :Lk], [p + head_off for p in attn_ptrs], Lp, (Lp * D, D, d),
                           variant=variant)
    px.barrier()  # every peer's head slices of my tokens have landed; peers are done reading their qkv_recv
    return attn_t[:n_attn].view(B * Lp, D)
