"""vc_time_embed_static: the timestep-embedding island for hosts that cannot allocate device memory.

Who needs it: a denoise loop whose device pool was handed out before the run and then closed, a plugin invoked on a
worker thread or callback that is not allowed near the allocator, a C host that itself received nothing but input and
output pointers.  Such a caller cannot supply the fp32 workspace `vc_time_embed` asks for, so the library keeps one of a
fixed size (328 KB per GPU) inside its own image and this entry point uses it.

What is checked here, through the C ABI (ctypes on libvc_b200.so, raw pointers):
  * two calls in a row, with different timesteps, each equal to `vc_time_embed` on the same inputs — the array is shared
    and still holds the first call's intermediates when the second starts, so the second call is the one that would
    show a stale-scratch bug.  Both entry points run the same four launches on the same arithmetic, so the tolerance is
    the tightest the suite uses: bit-identical;
  * no torch allocation happens inside the calls (outputs are made beforehand; the caching allocator's counters do
    not move), and a shape beyond the built-in limit is refused with a message instead of overrunning the array.

The file name sorts last on purpose: `pytest -x` stops at the first failure, so a failure here cannot hide the results of
the parity tests.  It does not depend on them either — it passes alone in a fresh process (then the first call below is
the one that looks up the array's device address) and after the rest of the suite."""
import ctypes
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_FLOATS = 8 * (256 + 2 * 5120)  # include/vc_b200.h: the documented limit of the built-in scratch


def _rand(shape, seed, scale):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(torch.bfloat16).cuda()


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _weights(D, freq_dim):
    return [_rand((D, freq_dim), 160, 0.02), _rand((D,), 161, 0.02), _rand((D, D), 162, 0.02), _rand((D,), 163, 0.02),
            _rand((6 * D, D), 164, 1 / math.sqrt(D)), _rand((6 * D,), 165, 0.02)]


@pytest.mark.parametrize("B,D,freq_dim", [(2, 1536, 256), (8, 5120, 256), (3, 256, 64)])
def test_static_scratch_matches_workspace_entry_point_twice(B, D, freq_dim):
    from versecrafter_b200 import _lib
    lib = _lib.load()
    w = _weights(D, freq_dim)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    # everything either entry point touches exists before the calls under test
    ts = [torch.linspace(999.0, 3.0, B, dtype=torch.float64, device="cuda"),
          torch.linspace(17.5, 640.0, B, dtype=torch.float64, device="cuda")]
    workspace = torch.empty(B * (freq_dim + 2 * D), dtype=torch.float32, device="cuda")
    e_ref = [torch.empty((B, D), dtype=torch.bfloat16, device="cuda") for _ in ts]
    e0_ref = [torch.empty((B, 6 * D), dtype=torch.bfloat16, device="cuda") for _ in ts]
    e_new = [torch.full((B, D), float("nan"), dtype=torch.bfloat16, device="cuda") for _ in ts]
    e0_new = [torch.full((B, 6 * D), float("nan"), dtype=torch.bfloat16, device="cuda") for _ in ts]
    for t, e, e0 in zip(ts, e_ref, e0_ref):
        _lib.check(lib.vc_time_embed(_p(t), B, freq_dim, D, *[_p(x) for x in w], _p(e), _p(e0), _p(workspace), stream),
                   "vc_time_embed")
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats()["allocation.all.allocated"]
    for t, e, e0 in zip(ts, e_new, e0_new):  # two calls in a row on the same stream: the second sees the first's leftovers
        _lib.check(lib.vc_time_embed_static(_p(t), B, freq_dim, D, *[_p(x) for x in w], _p(e), _p(e0), stream),
                   "vc_time_embed_static")
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats()["allocation.all.allocated"] == before
    for i in range(len(ts)):
        assert torch.isfinite(e_new[i].float()).all() and torch.isfinite(e0_new[i].float()).all()
        assert torch.equal(e_new[i], e_ref[i]), f"call {i}: e differs from vc_time_embed"
        assert torch.equal(e0_new[i], e0_ref[i]), f"call {i}: e0 differs from vc_time_embed"
    assert not torch.equal(e_ref[0], e_ref[1])  # the two calls really computed different things


def test_static_scratch_refuses_sizes_beyond_its_limit():
    from versecrafter_b200 import _lib
    lib = _lib.load()
    B, D, freq_dim = 8, 5128, 256  # 8 * (256 + 2*5128) = 84,096 > 83,968
    assert B * (freq_dim + 2 * D) > MAX_FLOATS
    dummy = torch.zeros(64, dtype=torch.float64, device="cuda")  # never dereferenced: the size check comes first
    rc = lib.vc_time_embed_static(_p(dummy), B, freq_dim, D, *[_p(dummy)] * 6, _p(dummy), _p(dummy),
                                  ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc < 0
    msg = lib.vc_last_error().decode()
    assert "exceed" in msg and str(MAX_FLOATS) in msg
    torch.cuda.synchronize()  # nothing was launched, nothing faulted
    with pytest.raises(_lib.VCError):
        _lib.check(rc, "vc_time_embed_static")
