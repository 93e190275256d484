"""N-GPU parity (needs >= 2 visible GPUs; skipped on a 1-GPU box): launches
tests/multigpu_parity.py under torchrun, one rank per GPU, and checks its verdict."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.parametrize("transport", ["peer", "nccl"])
@pytest.mark.parametrize("world", [2, 4])
def test_decomposed_run_equals_undecomposed_oracle(world, transport):
    """One process per GPU under torchrun, over the peer-memory transport (csrc/peer.cu: fused all-reduce, halo gate)
    and over the NCCL fallback (ncclSend/Recv + ncclAllReduce)."""
    if _ngpu() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(29600 + world + (10 if transport == "nccl" else 0)),
           os.path.join(HERE, "multigpu_parity.py")]
    env = dict(os.environ, CFD_B200_TRANSPORT=transport)
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env)
    assert out.returncode == 0 and "MULTIGPU PARITY OK" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]
    assert "transport=%s" % transport in out.stdout, out.stdout[-1500:]
