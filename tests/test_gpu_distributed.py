"""GPU, >= 2 devices: the multi-GPU plans end to end over NCCL (skipped on 1 GPU)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("exchange", ["split", "p2p", "fused", "nccl"])
def test_distributed_plans_over_nccl(exchange):
    """split: streaming split with bulk stores into peer memory (the default);
    p2p: copy engines into symmetric peer memory; fused: the partition kernel stores rows
    straight into the peers' buffers; nccl: all_to_all_single fallback."""
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    n = min(torch.cuda.device_count(), 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.join(HERE, "dist_gpu_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, DS_EXCHANGE=exchange))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert f"DIST_GPU_OK {n}" in r.stdout
