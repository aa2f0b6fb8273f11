"""rp_snapshot_async: the producer side of the reference's snapshot / trajectory recorders (Simulation.cpp:141-152,
RamsnapWriter.cpp:35-47) — the state of the moment of the call, delivered while later cycles run."""
import numpy as np
import pytest

import helpers as H
import rampack_b200 as rp
from rampack_b200 import NativeSimulation

pytestmark = pytest.mark.gpu


def _pinned(n):
    import torch
    return torch.empty((n, 12), dtype=torch.float64).pin_memory().numpy()


def test_async_snapshot_is_the_state_of_the_moment_it_was_requested():
    dev = rp.Device(0, seed=77)
    box, shapes = H.sc_lattice((28, 28, 16))
    dev.set_interaction(H.desc_sc())
    dev.upload(box, shapes)
    sim = NativeSimulation(dev, [(rp.RP_MOVE_ROTOTRANSLATION, 0.05, 0.05)], 1.0, 1.2, ("delta_triclinic", 0.001, False),
                           seed=77)
    sim.run(20)
    want_box, want = dev.download()
    want = want.copy()
    out = _pinned(dev.n)
    out[:] = -1.0
    got_box = dev.snapshot_async(out)
    sim.run(10)                       # queued behind the device-side copy: must not leak into the snapshot
    dev.snapshot_wait()
    assert dev.snapshot_done()
    assert np.array_equal(got_box, want_box)
    assert np.array_equal(out, want)
    _, after = dev.download()
    assert not np.array_equal(after, want)
    # a second request while the first may still be in flight: ordered on the device, both complete
    out2 = _pinned(dev.n)
    dev.snapshot_async(out)
    sim.run(2)
    dev.snapshot_async(out2)
    dev.snapshot_wait()
    _, now = dev.download()
    assert np.array_equal(out2, now)
    assert np.array_equal(out, after)
    sim.close()
    dev.close()


def test_snapshot_every_100_cycles_costs_under_one_per_cent():
    """`snapshot_every = 100` must not stall the sweep.  100 352 spherocylinders, 3 x 100 cycles with
    and without a snapshot after every block; device time by CUDA events."""
    import torch
    stream = torch.cuda.Stream()
    dev = rp.Device(0, seed=5, cuda_stream=stream.cuda_stream)
    box, shapes = H.sc_lattice((56, 56, 32))
    dev.set_interaction(H.desc_sc())
    dev.upload(box, shapes)
    sim = NativeSimulation(dev, [(rp.RP_MOVE_ROTOTRANSLATION, 0.05, 0.05)], 1.0, 1.2, None, seed=5)
    sim.run(50)
    out = _pinned(dev.n)
    dev.snapshot_async(out)           # (first call: allocates the device-side copy, creates the second stream)
    dev.snapshot_wait()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]

    def blocks(snap):
        for _ in range(3):
            sim.run(100)
            if snap:
                dev.snapshot_async(out)

    with torch.cuda.stream(stream):
        e[0].record(stream); blocks(False); e[1].record(stream)
        e[2].record(stream); blocks(True); e[3].record(stream)
    torch.cuda.synchronize()
    dev.snapshot_wait()
    plain, with_snap = e[0].elapsed_time(e[1]), e[2].elapsed_time(e[3])
    print(f"300 cycles: {plain:.2f} ms, with 3 async snapshots {with_snap:.2f} ms")
    assert with_snap < plain * 1.01 + 0.3
    sim.close()
    dev.close()
