"""Multi-GPU data plane on real peers (skipped below 2 GPUs): the NVLink peer-memory gradient exchanges (dense pull,
compacted pull, compacted push, owner-sliced push — the last is what bench.py --gpus N times — and the sharded Adam step
that consumes it) against the NCCL all-gather exchange on the same view
blocks AND against a plain NCCL all-reduce of every rank's ordinary single-view leaf gradients; all ranks must end
with bit-identical gradients. The 1-GPU box of the round-end test run skips this; bench.py's N>1 line carries the
same comparison as `grad_exchange_check`."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_peer_exchange_equals_nccl_on_real_peers(cuda_dev):
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (NVLink peers)")
    world = 2 if n < 4 else 4
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", "29541", os.path.join(ROOT, "tools", "peer_check.py"), "300000"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    try:                                  # the measured differences, kept with the run's other logs
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", f"peer_check_{world}gpu.log"), "w") as f:
            f.write(r.stdout)
    except OSError:
        pass
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-4000:]
    assert r.stdout.count("PEER_CHECK_OK") == world, r.stdout[-3000:]
