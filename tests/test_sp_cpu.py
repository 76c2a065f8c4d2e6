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
