"""Multi-GPU Ulysses parity (needs >= 2 visible GPUs; skipped on the single-GPU test box)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sp_equals_single_gpu(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(29500 + world), os.path.join(ROOT, "tools", "sp_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "OK" in r.stdout


def test_blocking_status_check_sees_a_timeout_at_once():
    """check_status(wait=True) must report a time-out recorded by work that is already enqueued (the last forward of a
    denoise loop), not only one from an earlier call; afterwards the exchange refuses further use.  One GPU: the status word
    is written the way the barrier kernel writes it."""
    from versecrafter_b200 import _lib, sp
    px = object.__new__(sp.PeerExchange)
    px.poisoned, px.timeout_ms, px.first_timeout_ms, px.first_epochs = None, 5000, 60000, 512
    px.barrier_impl, px._flags, px._epoch = "own", None, 0
    px._status = torch.zeros(1, dtype=torch.int32, device="cuda")
    px._status_host, px._status_event = None, None
    px.check_status()                 # first non-blocking call: queues a copy, nothing to look at yet
    px.check_status(wait=True)        # healthy
    px._status.fill_(-(1 << 31) | 5)  # 0x80000005, enqueued like a kernel of the current forward
    with pytest.raises(_lib.VCError, match="rank 5"):
        px.check_status(wait=True)
    with pytest.raises(_lib.VCError, match="unusable"):
        px.check_status()
    with pytest.raises(_lib.VCError, match="unusable"):
        px.barrier()
