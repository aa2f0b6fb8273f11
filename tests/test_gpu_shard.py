"""GPU test of the spatially sharded sweep (needs >= 2 GPUs; run with `gpurun --gpus 2`): an N-rank run over
NCCL must be BIT-IDENTICAL to the single-GPU run with the same seed — same counters, same box-move verdicts, same
final configuration on every rank."""
import os
import subprocess
import sys

import numpy as np
import pytest

import rampack_b200 as rp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def gpu_count():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("path", ["peer_memory", "nccl"])
@pytest.mark.parametrize("kind", [rp.RP_MOVE_ROTOTRANSLATION, rp.RP_MOVE_TRANSLATION])
def test_sharded_run_is_bit_identical_to_single_gpu(tmp_path, kind, path):
    """path: accepted moves written straight into the peers' replicas over NVLink peer memory (default), or the
    pack / ncclSend / ncclRecv / apply fallback (RP_SHARD_NCCL=1)."""
    n = gpu_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    world = 8 if n >= 8 else (4 if n >= 4 else 2)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    if path == "nccl":
        env["RP_SHARD_NCCL"] = "1"
    else:
        env.pop("RP_SHARD_NCCL", None)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                          "--master-addr", "127.0.0.1", "--master-port", "29541",
                          os.path.join(ROOT, "tests", "shard_worker.py"), str(tmp_path), str(kind)],
                         capture_output=True, text=True, timeout=900, env=env)
    print(out.stdout[-3000:])
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "RESULT ok=True" in out.stdout
    assert ("p2p=True" in out.stdout) == (path == "peer_memory"), "the peer-memory path did not come up"
    states = [np.load(tmp_path / f"state_{r}.npy") for r in range(world)]
    for r in range(1, world):
        assert np.array_equal(states[0], states[r]), f"rank {r} holds a different replica"
