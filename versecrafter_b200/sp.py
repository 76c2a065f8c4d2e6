"""Ulysses sequence parallelism around self-attention: every op of the DiT block is token-local
except self-attention, so each rank owns a contiguous chunk of Lp = L/P tokens of every sample (the reference's
torch.chunk(dim=1), wan_transformer3d_versecrafter.py:269-270, :366-367) and swaps "all heads of my tokens" for
"my heads of all tokens" with one all-to-all before attention and the inverse after it — what the reference gets
from xfuser's SeqAllToAll4D inside usp_attn_forward (not in the reference tree; ring attention is dropped by design).

The layout transforms are plain tensor permutations so they run on any backend: NCCL over NVLink on the B200 box,
gloo on CPU in tests/test_sp_cpu.py.
"""
import os

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
        self._side = None
        # schedule of ulysses_qkv_attention_overlapped (same arithmetic in every setting; tools/sp_overlap_ab.py measures them)
        self.overlap_mode = os.environ.get("VC_SP_OVERLAP_MODE", "full")   # full | v_only
        self.scatter_bulk = os.environ.get("VC_SP_SCATTER", "st") == "bulk"  # per-thread stores measured faster
        self.scatter_cap = int(os.environ.get("VC_SP_SCATTER_CAP", "0"))
        self.barrier_impl = os.environ.get("VC_SP_BARRIER", "own")         # own | torch
        self._flags, self._status, self._epoch = None, None, 0
        self._status_host, self._status_event = None, None
        # barrier time-out: VC_SP_BARRIER_TIMEOUT_MS for steady state; the first epochs of a process get at least
        # VC_SP_BARRIER_FIRST_TIMEOUT_MS because ranks reach their first forward seconds apart (module loading, first
        # rendezvous, a profiler or debugger attached to one rank)
        self.timeout_ms = int(os.environ.get("VC_SP_BARRIER_TIMEOUT_MS", "5000"))
        self.first_timeout_ms = max(self.timeout_ms, int(os.environ.get("VC_SP_BARRIER_FIRST_TIMEOUT_MS", "60000")))
        self.first_epochs = 512    # more than the barriers of one c3 forward (4 per block x 60 + GeoAdapter)
        self.poisoned = None       # message of the time-out that killed this group; every later call raises it

    def _timeout(self, epoch):
        return self.first_timeout_ms if epoch <= self.first_epochs else self.timeout_ms

    def _check_alive(self):
        from . import _lib
        if self.poisoned is not None:
            raise _lib.VCError(f"sequence-parallel group unusable: {self.poisoned} (the ranks are no longer in step; "
                               f"tear the process group down and start again)")

    def side_stream(self):
        """High-priority stream for the q / k scatter launches that run under the next GEMM (ulysses_qkv_attention_overlapped)."""
        if self._side is None:
            self._side = torch.cuda.Stream(device=self.device, priority=-1)
        return self._side

    def buffer(self, name, numel):
        """bf16 symmetric buffer of at least `numel` elements -> (tensor, [device address of every rank's copy])."""
        ent = self.bufs.get(name)
        if ent is None or ent[0].numel() < numel:
            t = self.symm.empty(int(numel), dtype=torch.bfloat16, device=self.device)
            h = self.symm.rendezvous(t, self.group.group_name)
            ent = (t, [int(p) for p in h.buffer_ptrs], h)
            self.bufs[name] = ent
        return ent[0], ent[1]

    def _ensure_flags(self):
        from . import ops
        if self._flags is None:
            t, ptrs = self.buffer("barrier_flags", 64)        # 128 bytes: 8 u32 epoch slots (+ padding) per rank
            t.zero_()
            self._status = torch.zeros(1, dtype=torch.int32, device=self.device)
            torch.cuda.current_stream().synchronize()
            self.bufs["barrier_flags"][2].barrier(channel=0)   # everybody's slots are zero before the first epoch
            torch.cuda.current_stream().synchronize()
            self._flags = ops._ptr_array(ptrs)

    def barrier(self):
        """Stream-ordered barrier of the group.  Default: this library's epoch-flag kernel (vc_peer_barrier, one NVLink
        round trip, csrc/peer_sync.cu); VC_SP_BARRIER=torch uses torch symmetric memory's barrier."""
        from . import _lib, ops
        with ops._timed("barrier", 0.0):
            if self.barrier_impl == "torch":
                next(iter(self.bufs.values()))[2].barrier(channel=0)
                return
            self._check_alive()
            self._ensure_flags()
            self._epoch += 1
            lib = _lib.load()
            _lib.check(lib.vc_peer_barrier(self._flags, self.rank, self.P, self._epoch & 0xFFFFFFFF, self._timeout(self._epoch),
                                           ops._ptr(self._status), ops._stream()), "vc_peer_barrier")
            ops.launch_count += 1

    def signal(self):
        """First half of a split barrier: tells every peer that everything enqueued so far on this stream (remote stores
        included) has completed.  Returns the epoch to hand to `wait`.  Every rank must call signal / wait / barrier in the
        same order (one epoch counter).  With VC_SP_BARRIER=torch there is no split form: signal is a no-op and wait is the
        full barrier."""
        from . import _lib, ops
        if self.barrier_impl == "torch":
            return None
        self._check_alive()
        self._ensure_flags()
        self._epoch += 1
        with ops._timed("barrier", 0.0):
            _lib.check(_lib.load().vc_peer_signal(self._flags, self.rank, self.P, self._epoch & 0xFFFFFFFF, ops._stream()),
                       "vc_peer_signal")
            ops.launch_count += 1
        return self._epoch

    def wait(self, epoch):
        """Second half: the stream continues once every peer has signalled `epoch` (or a later one)."""
        from . import _lib, ops
        if epoch is None:
            return self.barrier()
        self._check_alive()
        with ops._timed("barrier", 0.0):
            _lib.check(_lib.load().vc_peer_wait(self._flags, self.rank, self.P, epoch & 0xFFFFFFFF, self._timeout(epoch),
                                                ops._ptr(self._status), ops._stream()), "vc_peer_wait")
            ops.launch_count += 1

    def check_status(self, wait=False):
        """Raises if a barrier timed out (a peer never arrived).  wait=False: every call queues an asynchronous copy of the
        status word to pinned host memory and looks at the copy queued by the PREVIOUS call once its event has completed,
        so the launch queue is never drained — but a time-out inside the forward that has just been enqueued is only seen
        by the next call.  wait=True: copies the word after everything enqueued so far and synchronises on it, so it
        covers the current forward; the drop-in does this on the last step of a denoise loop (transformer.forward) and
        loop.denoise after its last step.  The status word on the GPU is sticky (every later wait returns at once, see
        csrc/peer_sync.cu), so a time-out poisons the group: this and every later barrier / signal / wait raise."""
        if self._status is None:
            return
        self._check_alive()
        if self._status_host is None:
            self._status_host = torch.zeros(1, dtype=torch.int32).pin_memory()
            self._status_event = torch.cuda.Event()
        elif self._status_event.query():
            self._raise_if_timed_out()
        elif not wait:
            return   # the previous copy is still in flight: look again next time
        self._status_host.copy_(self._status, non_blocking=True)
        self._status_event.record()
        if wait:
            self._status_event.synchronize()
            self._raise_if_timed_out()

    def _raise_if_timed_out(self):
        v = int(self._status_host[0])
        if v != 0:
            self.poisoned = f"barrier timed out waiting for rank {v & 0xff} (status {v & 0xffffffff:#x})"
            self._check_alive()


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
    ops.rmsnorm_rope_qkv_scatter(qkv, D, gain_q, gain_k, Lp, recv_ptrs, Hp, L, rope, grid, pos_offset=rank * Lp, eps=eps,
                                 bulk=px.scatter_bulk)
    px.barrier()  # every peer's q/k/v slices have landed here; peers are done reading attn_recv of the previous block
    r = recv_t[:n_recv].view(3, B, L, Hp, d)
    head_off = rank * Hp * d * 2  # bytes: this rank's heads inside a peer's [.., H, d] row
    Lk = L if k_true is None else min(k_true, L)  # padded tail tokens are not attended (k_lens, as on one GPU)
    ops.flash_attn_scatter(r[0], r[1][:, :Lk], r[2][:, :Lk], [p + head_off for p in attn_ptrs], Lp, (Lp * D, D, d),
                           variant=variant)
    px.barrier()  # every peer's head slices of my tokens have landed; peers are done reading their qkv_recv
    return attn_t[:n_attn].view(B * Lp, D)


def ulysses_fused_phases(px, qkv, B, Lp, H, d, gain_q, gain_k, rope, grid, eps, variant, k_true=None, tag=""):
    """ulysses_attention_fused cut at its two barriers, each barrier split into signal (end of the producing phase) and
    wait (start of the consuming phase), so that the caller can run another chain's kernels in between
    (transformer._run_blocks_interleaved).  Returns three callables; the last returns the attention output [B*Lp, H*d].
    `tag` selects a separate pair of symmetric buffers (two chains are in flight at once).  Same kernels, same arithmetic.
    Buffer reuse is ordered as in the un-split form: a chain's next scatter into the peers' qkv_recv follows its wait on the
    peers' post-attention signal; its next attention (writing the peers' attn_recv) follows its wait on the peers'
    post-scatter signal, which the peers issue after their previous o-projection has read attn_recv."""
    from . import ops
    P, rank = px.P, px.rank
    Hp, L, D = H // P, P * Lp, H * d
    n_recv, n_attn = 3 * B * L * Hp * d, B * Lp * D
    recv_t, recv_ptrs = px.buffer("qkv_recv" + tag, n_recv)
    attn_t, attn_ptrs = px.buffer("attn_recv" + tag, n_attn)
    ep = {}

    def scatter():
        ops.rmsnorm_rope_qkv_scatter(qkv, D, gain_q, gain_k, Lp, recv_ptrs, Hp, L, rope, grid, pos_offset=rank * Lp, eps=eps,
                                     bulk=px.scatter_bulk)
        ep[1] = px.signal()

    def attention():
        px.wait(ep[1])
        r = recv_t[:n_recv].view(3, B, L, Hp, d)
        head_off = rank * Hp * d * 2
        Lk = L if k_true is None else min(k_true, L)
        ops.flash_attn_scatter(r[0], r[1][:, :Lk], r[2][:, :Lk], [p + head_off for p in attn_ptrs], Lp, (Lp * D, D, d),
                               variant=variant)
        ep[2] = px.signal()

    def gathered():
        px.wait(ep[2])
        return attn_t[:n_attn].view(B * Lp, D)

    return scatter, attention, gathered


def ulysses_qkv_attention_overlapped(px, h, wqkv, bqkv, qkv, B, Lp, H, d, gain_q, gain_k, rope, grid, eps, variant,
                                     k_true=None):
    """h [B*Lp, D] (LayerNorm+modulate output) -> self-attention output [B*Lp, H*d] (view of the symmetric buffer), with the
    head-scatter all-to-alls hidden under the QKV projection instead of following it:
        main stream                              side stream (high priority)
        q = h Wq^T + bq          -> qkv[:, :D]
        k = h Wk^T + bk          -> qkv[:, D:2D]   RMSNorm+RoPE(q) -> peers' buffers   (vc_rmsnorm_rope_qkv_scatter, mask q)
        v = h Wv^T + bv -> peers' buffers          RMSNorm+RoPE(k) -> peers' buffers   (mask k)
          (vc_gemm_bf16_scatter: v needs no norm, its exchange is the GEMM's epilogue)
        barrier, attention (output rows stored to their owners), barrier
    The NVLink stores of q and k (two thirds of the exchanged bytes) run while the tensor cores work on the next projection
    and v never exists as a local matrix.  Arithmetic identical to ulysses_attention_fused (same kernels, same rounding)."""
    from . import ops
    P, rank = px.P, px.rank
    Hp, L, D = H // P, P * Lp, H * d
    n_recv, n_attn = 3 * B * L * Hp * d, B * Lp * D
    recv_t, recv_ptrs = px.buffer("qkv_recv", n_recv)
    attn_t, attn_ptrs = px.buffer("attn_recv", n_attn)
    main, side = torch.cuda.current_stream(), px.side_stream()
    scatter = dict(rows_per_batch=Lp, peer_ptrs=recv_ptrs, heads_per_peer=Hp, l_total=L, rope=rope, grid=grid,
                   pos_offset=rank * Lp, eps=eps, bulk=px.scatter_bulk, cta_cap=px.scatter_cap)
    v_block = 2 * B * L * Hp * d * 2  # bytes: offset of the v block inside a peer's [3, B, L, Hp, d] buffer
    v_ptrs = [p + v_block for p in recv_ptrs]
    if px.overlap_mode == "v_only":  # q|k as one GEMM and one scatter launch on this stream; only v rides in its GEMM
        ops.gemm(h, wqkv[:2 * D], bqkv[:2 * D], out=qkv[:, :2 * D])
        ops.gemm_scatter(h, wqkv[2 * D:], bqkv[2 * D:], Lp, v_ptrs, Hp, L, pos_offset=rank * Lp, head_dim=d)
        ops.rmsnorm_rope_qkv_scatter(qkv, D, gain_q, gain_k, which_mask=3, **scatter)
    else:
        ops.gemm(h, wqkv[:D], bqkv[:D], out=qkv[:, :D])
        ev_q = torch.cuda.Event()
        ev_q.record(main)
        ops.gemm(h, wqkv[D:2 * D], bqkv[D:2 * D], out=qkv[:, D:2 * D])
        ev_k = torch.cuda.Event()
        ev_k.record(main)
        with torch.cuda.stream(side):
            side.wait_event(ev_q)
            ops.rmsnorm_rope_qkv_scatter(qkv, D, gain_q, gain_k, which_mask=1, **scatter)
            side.wait_event(ev_k)
            ops.rmsnorm_rope_qkv_scatter(qkv, D, gain_q, gain_k, which_mask=2, **scatter)
            ev_s = torch.cuda.Event()
            ev_s.record(side)
        ops.gemm_scatter(h, wqkv[2 * D:], bqkv[2 * D:], Lp, v_ptrs, Hp, L, pos_offset=rank * Lp, head_dim=d)
        main.wait_event(ev_s)
    px.barrier()  # every peer's q/k/v slices have landed here; peers are done reading attn_recv of the previous block
    r = recv_t[:n_recv].view(3, B, L, Hp, d)
    head_off = rank * Hp * d * 2
    Lk = L if k_true is None else min(k_true, L)
    ops.flash_attn_scatter(r[0], r[1][:, :Lk], r[2][:, :Lk], [p + head_off for p in attn_ptrs], Lp, (Lp * D, D, d),
                           variant=variant)
    px.barrier()
    return attn_t[:n_attn].view(B * Lp, D)
