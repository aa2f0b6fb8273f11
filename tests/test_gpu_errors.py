"""Error behaviour of the C ABI on the device: the reference's Expects() preconditions (PreconditionException,
src/utils/Exceptions.h:19-66) surface as RP_ERR_PRECONDITION with the reference's condition text, and a rejected call
leaves the packing untouched."""
import numpy as np
import pytest

import rampack_b200 as rp
from rampack_b200 import capi
import helpers as H

pytestmark = pytest.mark.gpu


@pytest.fixture()
def device():
    dev = rp.Device(0, seed=99)
    yield dev
    dev.close()


def test_upload_rejects_shapes_outside_the_box_and_keeps_the_old_state(device):
    # Packing::reset: Expects(areShapesWithinBox(shapes, box)) (core/Packing.cpp:67-76, 1067-1073)
    device.set_interaction(H.desc_sc())
    box, shapes = H.sc_lattice((8, 8, 4))
    device.upload(box, shapes)
    bad = shapes.copy()
    bad[17, 0] = box[0, 0] + 0.25
    with pytest.raises(capi.PreconditionError, match="areShapesWithinBox"):
        device.upload(box, bad)
    device.n = shapes.shape[0]
    b, s = device.download()
    assert np.array_equal(s, shapes) and np.array_equal(b, box)
    bad = shapes.copy()
    bad[3, 2] = -1e-9
    with pytest.raises(capi.PreconditionError, match="areShapesWithinBox"):
        device.upload(box, bad)


def test_scaling_preconditions(device):
    # Packing::tryScaling: Expects(newBox.getVolume() != 0) (core/Packing.cpp:172-176); revertScaling needs a scaling
    device.set_interaction(H.desc_sc())
    box, shapes = H.sc_lattice((8, 8, 4))
    device.upload(box, shapes)
    with pytest.raises(capi.PreconditionError, match="revertScaling"):
        device.revert_scaling()
    flat = box.copy()
    flat[2, 2] = 0.0
    with pytest.raises(capi.PreconditionError, match="getVolume"):
        device.try_scaling(flat)
    assert device.try_scaling(box * 1.01) == 0.0
    device.revert_scaling()
    b, s = device.download()
    assert np.allclose(s, shapes, rtol=0, atol=1e-12)


def test_move_preconditions(device):
    # Packing::tryMove: Expects(particleIdx < size()) (core/Packing.cpp:84-90); acceptMove needs a tried move
    device.set_interaction(H.desc_sc())
    box, shapes = H.sc_lattice((8, 8, 4))
    device.upload(box, shapes)
    with pytest.raises(capi.PreconditionError, match="particleIdx"):
        device.try_move(shapes.shape[0], rp.RP_MOVE_TRANSLATION, (0.1, 0, 0))
    with pytest.raises(capi.PreconditionError, match="acceptMove"):
        device.accept_move()
    with pytest.raises(capi.PreconditionError):
        device.sweep(rp.RP_MOVE_TRANSLATION, 0.0, 0.1, 0)


def test_interaction_preconditions(device):
    with pytest.raises(capi.PreconditionError, match="radius"):
        device.set_interaction(rp.make_desc(hard=rp.RP_HARD_SPHERE, radius=0.0))
    with pytest.raises(capi.PreconditionError, match="epsilon"):
        device.set_interaction(rp.make_desc(soft=rp.RP_SOFT_LJ, epsilon=-1.0, sigma=1.0))
    with pytest.raises(capi.PreconditionError):
        device.set_interaction(rp.make_desc())


def test_cached_overlaps_needs_counting_mode(device):
    # Packing::getCachedNumberOfOverlaps throws std::runtime_error outside counting mode (core/Packing.cpp:1056-1060)
    device.set_interaction(H.desc_sc())
    box, shapes = H.sc_lattice((8, 8, 4))
    device.upload(box, shapes)
    with pytest.raises(capi.RpError):
        device.cached_overlaps()
    device.toggle_overlap_counting(True)
    assert device.cached_overlaps() == 0


def test_sweep_rejects_a_box_too_small_for_the_checkerboard(device):
    device.set_interaction(H.desc_sc())
    box = np.diag([7.0, 30.0, 30.0])
    shapes = np.zeros((1, 12))
    shapes[0, :3] = (3.5, 15, 15)
    shapes[0, 3:] = np.eye(3).ravel()
    device.upload(box, shapes)
    with pytest.raises(capi.RpError, match="checkerboard"):
        device.sweep(rp.RP_MOVE_TRANSLATION, 0.1, 0.1, 0)


def test_renormalize_orientations(device):
    """Packing::renormalizeOrientations / fixRotationMatrix (core/Packing.cpp:1178-1219): three iterations of
    R <- 1.5 R - 0.5 R R^T R restore orthogonality; positions stay; no overlap is created."""
    device.set_interaction(H.desc_sc())
    box, shapes = H.sc_lattice((8, 8, 4))
    rng = np.random.default_rng(3)
    drift = shapes.copy()
    drift[:, 3:] += rng.normal(scale=1e-6, size=(shapes.shape[0], 9))        # accumulated round-off, exaggerated
    device.upload(box, drift)
    assert device.renormalize_orientations() == 0
    _, s = device.download()
    want = drift[:, 3:].reshape(-1, 3, 3).copy()
    for _ in range(3):
        want = 1.5 * want - (0.5 * want) @ np.transpose(want, (0, 2, 1)) @ want
    got = s[:, 3:].reshape(-1, 3, 3)
    assert np.allclose(got, want, rtol=0, atol=4e-15)
    assert np.abs(got @ np.transpose(got, (0, 2, 1)) - np.eye(3)).max() < 2e-15
    assert np.array_equal(s[:, :3], drift[:, :3])
    assert device.count_total_overlaps(False) == 0
    # counting mode: the reference discards the fixed matrices (it resets from a copy taken before the fix)
    device.upload(box, drift)
    device.toggle_overlap_counting(True)
    assert device.renormalize_orientations(allow_overlaps=True) == 0
    _, s2 = device.download()
    assert np.array_equal(s2, drift)
