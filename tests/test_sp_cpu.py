"""Host-side logic of the Ulysses path on CPU: 2 processes over gloo.  Checks that scatter_heads / gather_heads are
the all-to-all permutations the sequence-parallel attention needs (equal to slicing the un-sharded tensor) and that
an attention computed on [all tokens, my heads] and gathered back equals the single-process result."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import restate
        from versecrafter_b200 import sp
        B, L, H, d = 2, 12, 4, 8
        g = torch.Generator().manual_seed(0)
        q, k, v = (torch.randn(B, L, H, d, generator=g) for _ in range(3))
        Lp, Hp = L // world, H // world
        mine = slice(rank * Lp, (rank + 1) * Lp)
        heads = slice(rank * Hp, (rank + 1) * Hp)
        qs = sp.scatter_heads(q[:, mine].contiguous(), world)
        assert torch.equal(qs, q[:, :, heads]), "scatter_heads != slicing"
        ks, vs = sp.scatter_heads(k[:, mine].contiguous(), world), sp.scatter_heads(v[:, mine].contiguous(), world)
        o_local = restate.sdpa(qs, ks, vs)                      # my heads, all tokens
        o = sp.gather_heads(o_local, world)                     # all heads, my tokens
        want = restate.sdpa(q, k, v)[:, mine]
        assert torch.allclose(o, want, atol=1e-5), "ulysses attention != single-process attention"
        full = sp.all_gather_tokens(want.contiguous(), 1)
        assert torch.allclose(full, restate.sdpa(q, k, v), atol=1e-5)
        ret[rank] = "ok"
    except Exception as e:  # noqa: BLE001
        ret[rank] = repr(e)
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_ulysses_layouts_world2():
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, port, ret), nprocs=2, join=True)
    assert dict(ret) == {0: "ok", 1: "ok"}, dict(ret)


def test_fused_scatter_layouts_equal_all_to_all_slices():
    """The destination formulas of the fused kernels (sp.recv_layout_index / attn_layout_index, mirrored in
    norm_kernels.cu / flash_attn_tcgen05.cu) place every element exactly where scatter_heads / gather_heads would."""
    from versecrafter_b200 import sp
    B, L, H, d, P = 2, 12, 8, 4, 4
    Lp, Hp = L // P, H // P
    g = torch.Generator().manual_seed(1)
    qkv = torch.randn(3, B, L, H, d, generator=g)
    recv = [torch.full((3 * B * L * Hp * d,), float("nan")) for _ in range(P)]
    for w in range(3):
        for b in range(B):
            for t in range(L):
                for h in range(H):
                    peer, off = sp.recv_layout_index(w, b, t, h, B, L, Hp, d)
                    recv[peer][off:off + d] = qkv[w, b, t, h]
    for p in range(P):
        assert torch.equal(recv[p].view(3, B, L, Hp, d), qkv[:, :, :, p * Hp:(p + 1) * Hp])
    o = torch.randn(B, L, H, d, generator=g)   # attention output, produced per rank as [B, L, Hp, d]
    attn = [torch.full((B * Lp * H * d,), float("nan")) for _ in range(P)]
    for b in range(B):
        for t in range(L):
            for h in range(H):
                peer, off = sp.attn_layout_index(b, t, h, Lp, H, d)
                attn[peer][off:off + d] = o[b, t, h]
    for r in range(P):
        assert torch.equal(attn[r].view(B, Lp, H, d), o[:, r * Lp:(r + 1) * Lp])


def test_bench_work_formula_matches_survey():
    import bench
    cfg, grid, _ = bench.CONFIGS["c3"]
    gemm, attn = bench.work_per_step(cfg, 32760, 2)
    assert abs(gemm / 1e15 - 2.429) < 0.01 and abs(attn / 1e15 - 2.679) < 0.01


@pytest.mark.parametrize("num_layers,geo_layers", [(8, None), (6, [0, 1, 4]), (5, [0, 3])])
def test_interleaved_schedule_order(monkeypatch, num_layers, geo_layers):
    """Host logic of transformer._run_blocks_interleaved (no GPU: the phases and the GEMM are replaced by recorders).
    Checks the order it enqueues: each chain's phases in order; phase [2] of GeoAdapter block j+1 (which updates c in
    place) after `after_proj(c_j)` has read c; hint j consumed right after backbone layer geoada_layers[j], with block j
    complete; a backbone WAIT ([1], [2]) is preceded by a control phase whenever the control chain still has one that may run."""
    from types import SimpleNamespace
    from versecrafter_b200 import ops, transformer
    cfg = dict(dim=128, ffn_dim=256, num_heads=1, num_layers=num_layers, text_dim=32, geoada_in_dim=128)
    if geo_layers is not None:
        cfg["geoada_layers"] = geo_layers
    model = transformer.VerseCrafterWanTransformer3DModel(**cfg, device="cpu", dtype=torch.bfloat16, init=False)
    log = []
    blocks = {id(b): ("x", i) for i, b in enumerate(model.blocks)}
    blocks.update({id(b): ("c", j) for j, b in enumerate(model.geoada_blocks)})

    def fake_phases(blk, x2d, m, ctx, st, tag=""):
        chain, idx = blocks[id(blk)]
        assert tag == ("" if chain == "x" else ".c")
        return [lambda k=k: log.append((chain, idx, k)) for k in range(4)]

    def fake_gemm(a, w, b=None, **kw):
        owner = next(("before", j) if (hasattr(g, "before_proj") and w is g.before_proj.weight) else ("after", j)
                     for j, g in enumerate(model.geoada_blocks)
                     if (hasattr(g, "before_proj") and w is g.before_proj.weight) or w is g.after_proj.weight)
        log.append(owner + (None,))
        return a

    monkeypatch.setattr(model, "_block_phases", fake_phases)
    monkeypatch.setattr(ops, "gemm", fake_gemm)
    n_geo = len(model.geoada_blocks)
    st = SimpleNamespace(B=1, Lp=4, m_all=[None] * (num_layers + n_geo), c_emb=torch.zeros(4, 128))
    x2d, c2d = torch.zeros(4, 128), torch.zeros(4, 128)
    model._run_blocks_interleaved(x2d, c2d, None, None, st, 1.0)

    pos = {e: n for n, e in enumerate(log)}
    assert len(pos) == len(log)                                            # nothing enqueued twice
    for chain, n in (("x", num_layers), ("c", n_geo)):
        seq = [e for e in log if e[0] == chain]
        assert seq == [(chain, i, k) for i in range(n) for k in range(4)]   # every phase once, in chain order
    layers = model.geoada_layers
    assert pos[("before", 0, None)] < pos[("c", 0, 0)] and pos[("before", 0, None)] < pos[("x", 0, 2)]   # reads x0
    for j in range(n_geo):
        a = pos[("after", j, None)]
        assert pos[("c", j, 3)] < a and pos[("x", layers[j], 3)] < a        # block j and its layer are complete
        if layers[j] + 1 < num_layers:
            assert a < pos[("x", layers[j] + 1, 0)]                         # the hint is in x before the next layer reads it
        if j + 1 < n_geo:
            assert a < pos[("c", j + 1, 2)]                                 # c_j read before block j+1 overwrites c
    # slack: between a backbone SIGNAL (end of [0] / [1]) and its WAIT (start of [1] / [2]) sits a control phase, as long
    # as the control chain has one whose constraints allow it
    covered = sum(1 for i in range(num_layers) for k in (0, 1) if log[pos[("x", i, k)] + 1][0] == "c")
    if geo_layers is None:   # every second layer: all control phases sit in such a slot except the two drained at the end of
        assert covered == 4 * n_geo - 3   # layer 0 (hint 0 is due there); block 0's first slot starts with before_proj
    else:
        assert covered >= n_geo


def _bare_exchange():
    """A PeerExchange without a process group: only the time-out bookkeeping (the fields check_status touches)."""
    from versecrafter_b200 import sp
    px = object.__new__(sp.PeerExchange)
    px.poisoned, px.timeout_ms, px.first_timeout_ms, px.first_epochs = None, 5000, 60000, 512
    px.barrier_impl, px._flags, px._epoch = "own", None, 0
    px._status, px._status_host, px._status_event = None, None, None
    return px


def test_barrier_timeout_poisons_the_exchange():
    """A barrier that timed out must not be survivable: the status word on the GPU is sticky (later waits return at once), so
    after the first report every barrier / signal / wait / check raises instead of running unsynchronised."""
    from versecrafter_b200 import _lib
    px = _bare_exchange()
    px._check_alive()                                            # healthy: no exception
    px._status_host = torch.tensor([0], dtype=torch.int32)
    px._raise_if_timed_out()                                     # status 0: nothing
    px._status_host = torch.tensor([-(1 << 31) | 3], dtype=torch.int32)   # 0x80000003: rank 3 never arrived
    with pytest.raises(_lib.VCError, match="rank 3"):
        px._raise_if_timed_out()
    assert px.poisoned is not None
    for call in (px.barrier, px.signal, lambda: px.wait(7), px._check_alive):
        with pytest.raises(_lib.VCError, match="unusable"):
            call()


def test_barrier_timeout_is_longer_for_the_first_epochs(monkeypatch):
    px = _bare_exchange()
    assert px._timeout(1) == 60000 and px._timeout(512) == 60000 and px._timeout(513) == 5000
    px.timeout_ms, px.first_timeout_ms = 90000, 90000            # VC_SP_BARRIER_TIMEOUT_MS above the first-epoch default
    assert px._timeout(1) == 90000 and px._timeout(10 ** 6) == 90000
