"""CPU tests of the sharded sweep's host logic (no device): slab partition and halo plan, alone and across a
world_size-2 gloo group (the plumbing bench.py uses to distribute the NCCL id)."""
import os
import subprocess
import sys

import pytest

from rampack_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("nz,P", [(4, 2), (12, 2), (74, 8), (116, 8), (16, 8), (10, 3)])
def test_slabs_tile_the_layers_and_plans_match(nz, P):
    plans = [[capi.shard_plan(nz, P, r, c) for c in range(8)] for r in range(P)]
    # slabs: contiguous, non-empty, cover all active layers
    assert plans[0][0]["a_lo"] == 0 and plans[P - 1][0]["a_hi"] == nz // 2
    for r in range(P):
        assert plans[r][0]["a_hi"] > plans[r][0]["a_lo"]
        if r:
            assert plans[r][0]["a_lo"] == plans[r - 1][0]["a_hi"]
    for c in range(8):
        for r in range(P):
            p = plans[r][c]
            q = plans[p["peer_send"]][c]
            # what I send is exactly what my peer expects to receive, and it comes from me
            assert q["recv_layer"] == p["send_layer"] and q["peer_recv"] == r
            lo, hi = 2 * p["a_lo"], 2 * p["a_hi"]
            assert lo <= p["send_layer"] < hi                       # I only send my own cells
            assert not (lo <= p["recv_layer"] < hi) or P == 1       # and receive foreign ones
            # the layer sent has the parity of the active cells; the receiver's adjacent layer has the other one
            assert p["send_layer"] % 2 == (c >> 2) & 1
            # the received layer touches my slab (periodic)
            assert (p["recv_layer"] - hi) % nz == 0 or (lo - 1 - p["recv_layer"]) % nz == 0


def test_plan_rejects_too_many_ranks():
    with pytest.raises(capi.RpError):
        capi.shard_plan(6, 4, 0, 0)
    with pytest.raises(capi.RpError):
        capi.shard_plan(7, 2, 0, 0)


WORKER = r"""
import os, sys, json
sys.path.insert(0, sys.argv[1])
import torch, torch.distributed as dist
from rampack_b200 import capi
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
# the id plumbing of bench.py: rank 0 makes 128 bytes, everybody receives them
ident = torch.arange(128, dtype=torch.uint8) if rank == 0 else torch.zeros(128, dtype=torch.uint8)
dist.broadcast(ident, 0)
assert ident.tolist() == list(range(128))
nz = 20
mine = [capi.shard_plan(nz, world, rank, c) for c in range(8)]
t = torch.tensor([[p["send_layer"], p["recv_layer"], p["peer_send"], p["peer_recv"]] for p in mine])
allp = [torch.zeros_like(t) for _ in range(world)]
dist.all_gather(allp, t)
for c in range(8):
    send_layer, recv_layer, peer_send, peer_recv = t[c].tolist()
    assert allp[peer_send][c][1].item() == send_layer and allp[peer_send][c][3].item() == rank
    assert allp[peer_recv][c][0].item() == recv_layer and allp[peer_recv][c][2].item() == rank
dist.barrier()
if rank == 0:
    print("OK")
dist.destroy_process_group()
"""


def test_halo_plan_consistent_across_a_gloo_group(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29531", str(script), ROOT],
                         capture_output=True, text=True, timeout=300, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "OK" in out.stdout
