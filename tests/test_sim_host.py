"""Host logic of rp_sim (no device): the box proposals of the reference's scalers
(src/core/volume_scalers/DeltaVolumeScaler.cpp, Linear/LogScalingFactorSampler.cpp, AnisotropicVolumeScaler.cpp:18-33,
TriclinicDeltaScaler.cpp:7-25, TriclinicAdapter.h:27-33) and the configuration preconditions."""
import os

import numpy as np
import pytest

import rampack_b200 as rp

BOX = np.array([[10.0, 1.0, 0.5], [0.0, 12.0, 0.7], [0.0, 0.0, 15.0]])
ROTO = [(rp.RP_MOVE_ROTOTRANSLATION, 0.1, 0.1)]


def heights(box):
    s1, s2, s3 = box[:, 0], box[:, 1], box[:, 2]
    v = abs(np.dot(np.cross(s1, s2), s3))
    return np.array([v / np.linalg.norm(np.cross(s2, s3)), v / np.linalg.norm(np.cross(s3, s1)),
                     v / np.linalg.norm(np.cross(s1, s2))])


def column_factors(new, old):
    f = []
    for c in range(3):
        ratio = new[:, c][old[:, c] != 0] / old[:, c][old[:, c] != 0]
        assert np.allclose(ratio, ratio[0], rtol=1e-14)
        f.append(ratio[0])
    return np.array(f)


def test_delta_v_scales_all_sides_by_the_cube_root_of_the_height_volume_ratio():
    sim = rp.NativeSimulation(None, ROTO, scaler=("delta_v", 30.0), seed=1)
    hv = np.prod(heights(BOX))
    seen = []
    for _ in range(200):
        f = column_factors(sim.propose_box(BOX), BOX)
        assert np.allclose(f, f[0], rtol=1e-15)
        dv = (f[0] ** 3 - 1) * hv
        assert abs(dv) <= 30.0 * (1 + 1e-12)
        seen.append(dv)
    assert min(seen) < -20 and max(seen) > 20


@pytest.mark.parametrize("kind", ["linear", "log"])
def test_anisotropic_independent_changes_exactly_one_group(kind):
    sim = rp.NativeSimulation(None, ROTO, scaler=(kind, 0.2, [3, 4], True), seed=2)
    groups = set()
    h = heights(BOX)
    for _ in range(200):
        f = column_factors(sim.propose_box(BOX), BOX)
        changed = tuple(np.abs(f - 1) > 1e-15)
        assert changed in ((True, True, False), (False, False, True))
        groups.add(changed)
        if kind == "log":
            assert np.all(np.abs(np.log(f)) <= 0.2 * (1 + 1e-12))
            if changed[0]:
                assert abs(f[0] - f[1]) < 1e-14
        else:
            step = (f - 1) * h
            assert np.all(np.abs(step) <= 0.2 * (1 + 1e-9))
            if changed[0]:
                assert abs(step[0] - step[1]) < 1e-9          # one sampled step for the whole group
    assert len(groups) == 2


def test_anisotropic_dependent_multiplies_the_groups():
    sim = rp.NativeSimulation(None, ROTO, scaler=("log", 0.1, [1, 2, 4], False), seed=3)
    f = np.array([column_factors(sim.propose_box(BOX), BOX) for _ in range(100)])
    assert np.all(np.abs(np.log(f)) <= 0.1 * (1 + 1e-12))
    assert np.all(np.abs(f[:, 0] - f[:, 1]) > 0)                # three independent draws


@pytest.mark.parametrize("independent", [True, False])
def test_triclinic_delta(independent):
    sim = rp.NativeSimulation(None, ROTO, scaler=("delta_triclinic", 0.05, independent), seed=4)
    sides = set()
    for _ in range(100):
        d = sim.propose_box(BOX) - BOX
        assert np.all(np.abs(d) <= 0.05)
        cols = tuple(np.any(d != 0, axis=0))
        if independent:
            assert sum(cols) == 1
            sides.add(cols)
        else:
            assert np.count_nonzero(d) == 9
    assert not independent or len(sides) == 3


def test_config_preconditions():
    with pytest.raises(rp.PreconditionError):
        rp.NativeSimulation(None, [(rp.RP_MOVE_TRANSLATION, 0.0, 0.1)])
    with pytest.raises(rp.PreconditionError):
        rp.NativeSimulation(None, ROTO, temperature=0.0)
    with pytest.raises(rp.PreconditionError):
        rp.NativeSimulation(None, ROTO, scaler=("delta_v", -1.0))
    with pytest.raises(rp.PreconditionError):
        rp.NativeSimulation(None, ROTO, scaler=("linear", 0.1, [0]))
    sim = rp.NativeSimulation(None, ROTO)           # NVT: no scaler to ask
    with pytest.raises(rp.RpError):
        sim.propose_box(BOX)


# ---- the box proposals against the reference's own scaler classes (VERDICT W4) -------------------------------------
def _ref_available():
    from oracle import refwrap as R
    return os.path.exists(R.ref_path())


SCALER_CASES = [
    ("delta_v", ("delta_v", 3.0), dict(kind="delta_v", step=3.0, independent=False, groups=[])),
    ("linear_xyz", ("linear", 0.2, [7], False), dict(kind="linear", step=0.2, independent=False, groups=[7])),
    ("linear_x_yz", ("linear", 0.2, [1, 6], False), dict(kind="linear", step=0.2, independent=False, groups=[1, 6])),
    ("linear_x_y_z_ind", ("linear", 0.3, [1, 2, 4], True),
     dict(kind="linear", step=0.3, independent=True, groups=[1, 2, 4])),
    ("log_xy_z", ("log", 0.05, [3, 4], False), dict(kind="log", step=0.05, independent=False, groups=[3, 4])),
    ("log_x_y_z_ind", ("log", 0.05, [1, 2, 4], True), dict(kind="log", step=0.05, independent=True, groups=[1, 2, 4])),
    ("triclinic_together", ("delta_triclinic", 0.01, False),
     dict(kind="delta_triclinic", step=0.01, independent=False, groups=[])),
    ("triclinic_independent", ("delta_triclinic", 0.01, True),
     dict(kind="delta_triclinic", step=0.01, independent=True, groups=[])),
]


@pytest.mark.parametrize("name,scaler,ref", SCALER_CASES, ids=[c[0] for c in SCALER_CASES])
def test_box_proposals_equal_the_reference_scaler_classes(name, scaler, ref):
    """rp_sim_propose_box draws from std::mt19937 with the reference's distributions in the reference's order, so a
    sequence of proposals from the same seed must equal TriclinicBoxScaler::updateBox of the reference's classes
    (core/volume_scalers/*.cpp, TriclinicAdapter.h:27-33) to the last bit (cbrt / exp: the same libm on both sides)."""
    if not _ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    from oracle import refwrap as R
    box = np.array([[10.0, 0.5, -0.3], [0.2, 12.0, 0.7], [-0.4, 0.1, 9.0]])
    seed, count = 20261, 300
    sim = rp.NativeSimulation(None, [(rp.RP_MOVE_ROTOTRANSLATION, 0.1, 0.1)], 1.0, 1.0, scaler, seed=seed)
    got = np.array([sim.propose_box(box) for _ in range(count)])
    want = R.ref_propose_boxes(seed=seed, old_box=box, count=count, **ref)
    assert np.array_equal(got, want), np.abs(got - want).max()
    sim.close()
