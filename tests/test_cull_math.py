"""The second FP32 cull of the sweep kernel (segmentCullPass, rampack_b200/csrc/rp_sweep_kernel.cuh) restated in numpy
float32: a pair it drops must be disjoint for the oracle's exact spherocylinder test (SpherocylinderTraits.cpp:28-46
with SegmentDistanceCalculator.h:22-85), and its squared distance must stay far inside the margin the device uses
(1e-3 * range^2).  CPU test of the formula; the device kernel itself is checked against the reference on the GPU
(test_gpu_sweep_replay.py, test_sweep_fp32_cull_never_changes_a_verdict)."""
import numpy as np
import pytest

import rampack_b200 as rp
from oracle import refwrap as R

L, RADIUS = 5.0, 0.5
HALF = 0.5 * L
F = np.float32


def cull_distance2(c1, h1, c2, h2, origin):
    """float32 restatement of segmentCullPass: inputs are rounded the way the kernel rounds them (positions relative
    to the cell centre, trial axis as a unit vector times the half length, candidate half axis directly)."""
    p1 = (c1 - origin).astype(F)
    p2 = (c2 - origin).astype(F)
    d = p1 - p2
    h1 = (h1 / HALF).astype(F) * F(HALF)
    h2 = h2.astype(F)
    a = F(HALF) * F(HALF)
    b = (h1 * h2).sum(1, dtype=F)
    cc = (h1 * d).sum(1, dtype=F)
    f = (h2 * d).sum(1, dtype=F)
    denom = a * a - b * b
    with np.errstate(all="ignore"):
        s = np.clip(((b * f - cc * a) / np.maximum(denom, F(1e-20))).astype(F), -1, 1).astype(F)
    t0 = ((b * s + f) / a).astype(F)
    t = np.clip(t0, -1, 1).astype(F)
    s1 = np.clip(((b * t - cc) / a).astype(F), -1, 1).astype(F)
    s = np.where(t != t0, s1, s)
    p = d + s[:, None] * h1 - t[:, None] * h2
    return (p * p).sum(1, dtype=F).astype(np.float64)


def brute_distance2(c1, h1, c2, h2, steps=801):
    """true minimum to ~1e-5: scan s, the best t for every s in closed form"""
    best = np.full(c1.shape[0], np.inf)
    for s in np.linspace(-1, 1, steps):
        p0 = (c1 - c2) + s * h1
        t = np.clip((p0 * h2).sum(1) / (h2 * h2).sum(1), -1, 1)
        p = p0 - t[:, None] * h2
        best = np.minimum(best, (p * p).sum(1))
    return best


def make_pairs(rng, n, mode):
    u1 = rng.normal(size=(n, 3))
    u1 /= np.linalg.norm(u1, axis=1)[:, None]
    if mode == "random":
        u2 = rng.normal(size=(n, 3))
    elif mode == "parallel":
        u2 = u1 + rng.normal(size=(n, 3)) * 10.0 ** rng.uniform(-9, -2, size=(n, 1))
    else:
        u2 = -u1 + rng.normal(size=(n, 3)) * 10.0 ** rng.uniform(-9, -2, size=(n, 1))
    u2 /= np.linalg.norm(u2, axis=1)[:, None]
    origin = rng.uniform(-50, 50, size=(1, 3))                  # the cell centre, anywhere in a large box
    c1 = origin + rng.uniform(-5, 5, size=(n, 3))
    dirn = rng.normal(size=(n, 3))
    dirn /= np.linalg.norm(dirn, axis=1)[:, None]
    c2 = c1 + dirn * rng.uniform(0.0, 4.0, size=(n, 1))
    return c1, u1, c2, u2, origin


def frames(u, rng):
    """orientation matrices whose third column (the spherocylinder axis) is u"""
    a = rng.normal(size=u.shape)
    a -= (a * u).sum(1)[:, None] * u
    a /= np.linalg.norm(a, axis=1)[:, None]
    b = np.cross(u, a)
    return np.stack([a, b, u], axis=2).reshape(-1, 9)


@pytest.mark.parametrize("mode", ["random", "parallel", "antiparallel"])
def test_segment_cull_never_drops_an_overlapping_pair(mode):
    rng = np.random.default_rng({"random": 1, "parallel": 2, "antiparallel": 3}[mode])
    n = 400_000
    c1, u1, c2, u2, origin = make_pairs(rng, n, mode)
    pairs = np.zeros(n, dtype=rp.PAIR_DTYPE)
    pairs["pos1"], pairs["pos2"] = c1, c2
    pairs["orientation1"], pairs["orientation2"] = frames(u1, rng), frames(u2, rng)
    # the half axes as the device records hold them: orientation column 2 times l / 2
    h1 = pairs["orientation1"][:, [2, 5, 8]] * HALF
    h2 = pairs["orientation2"][:, [2, 5, 8]] * HALF
    desc = rp.make_desc(hard=rp.RP_HARD_SPHEROCYLINDER, radius=RADIUS, length=L)
    traits = R.RefTraits(desc) if R.have_ref() else R.PortTraits(desc)
    verdict, _ = traits.pair_test_batch(pairs)
    thr2 = (2 * RADIUS) ** 2 + 1e-3 * (L + 2 * RADIUS) ** 2
    d2 = cull_distance2(c1, h1, c2, h2, origin)
    dropped = d2 >= thr2
    assert verdict.sum() > n // 50 and dropped.sum() > n // 4            # both outcomes are well populated
    assert not np.any(dropped & (verdict != 0))
    # the margin is not what saves it: near contact the FP32 value is within 1e-4 of the true minimum
    near = np.nonzero(d2 < 1.5)[0][:20000]
    excess = d2[near] - brute_distance2(c1[near], h1[near], c2[near], h2[near])
    assert excess.max() < 1e-4, excess.max()
