"""north_star parity run: >= 10^8 spherocylinder pair tests drawn from snapshots written by the REFERENCE's own
Simulation, device verdicts (rp_pair_test_batch) against the reference's overlapBetween — bit-exact — with every pair
within 1e-12 relative of contact listed explicitly.  Test infrastructure (uses oracle/_ref); run on the GPU box:

    python tools/pairs_1e8.py [--target 1e8] [--workload sc_1m] > gpurun_out/pairs_1e8.json

Snapshots: the workload lattice, equilibrated on the device (NVT, `--equil` cycles), then `--ref-cycles` NVT cycles of
the reference's Simulation::integrate on the host -> snapshot; repeated until the target is reached.  Pairs: all
unordered pairs (i, j) whose mass centres are closer than getRangeRadius() = l + 2r in the minimum image (the pairs
the reference's pre-filter lets through, SpherocylinderTraits.cpp:34-36), as explicit rp_pair records with the lattice
translation the reference's HardcodedTranslation would carry; particle i carries a random trial rototranslation
(translation +-0.5, rotation up to 0.3 rad) — the pair test a Monte Carlo trial performs — so that a few per cent of
the pairs overlap (in a valid hard-particle snapshot every untouched pair is disjoint).
"""
import argparse
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import rampack_b200 as rp  # noqa: E402
from rampack_b200 import workloads as W  # noqa: E402
from oracle import refwrap as R  # noqa: E402


def seg_dist2(p1, h1, p2, h2):
    """Squared distance between the segments p1 +- h1 and p2 +- h2 (numpy restatement of
    SegmentDistanceCalculator.h:22-85; only used to LIST near-contact pairs, never to decide a verdict)."""
    a0, b0 = p1 - h1, p2 - h2
    u, v, w = 2 * h1, 2 * h2, a0 - b0
    a, b, c = (u * u).sum(1), (u * v).sum(1), (v * v).sum(1)
    d, e = (u * w).sum(1), (v * w).sum(1)
    D = a * c - b * b
    eps = 1e-10
    par = D < eps
    sN = np.where(par, 0.0, b * e - c * d)
    sD = np.where(par, 1.0, D)
    tN = np.where(par, e, a * e - b * d)
    tD = np.where(par, c, D)
    lo = (~par) & (sN < 0)
    sN = np.where(lo, 0.0, sN); tN = np.where(lo, e, tN); tD = np.where(lo, c, tD)
    hi = (~par) & (~lo) & (sN > sD)
    sN = np.where(hi, sD, sN); tN = np.where(hi, e + b, tN); tD = np.where(hi, c, tD)
    tneg = tN < 0
    tbig = (~tneg) & (tN > tD)
    sN2 = np.where(tneg, np.where(-d < 0, 0.0, np.where(-d > a, sD, -d)),
                   np.where(tbig, np.where(-d + b < 0, 0.0, np.where(-d + b > a, sD, -d + b)), sN))
    sD2 = np.where((tneg & (-d >= 0) & (-d <= a)) | (tbig & (-d + b >= 0) & (-d + b <= a)), a, sD)
    tN = np.where(tneg, 0.0, np.where(tbig, tD, tN))
    with np.errstate(divide="ignore", invalid="ignore"):
        sc = np.where(np.abs(sN2) < eps, 0.0, sN2 / sD2)
        tc = np.where(np.abs(tN) < eps, 0.0, tN / tD)
    dp = w + sc[:, None] * u - tc[:, None] * v
    return (dp * dp).sum(1)


def neighbour_pairs(box, shapes, cutoff):
    from scipy.spatial import cKDTree
    L = np.array([box[0, 0], box[1, 1], box[2, 2]])
    pos = np.mod(shapes[:, :3], L)
    tree = cKDTree(pos, boxsize=L)
    return tree.query_pairs(cutoff, output_type="ndarray")


def build_pairs(box, shapes, ij, rng, trans_step=0.5, rot_step=0.3):
    """Particle i takes a random trial rototranslation (what the sweep tests against its neighbours: in a valid
    hard-particle snapshot every untouched pair is disjoint), particle j stays as the snapshot has it."""
    L = np.array([box[0, 0], box[1, 1], box[2, 2]])
    i, j = ij[:, 0], ij[:, 1]
    n = ij.shape[0]
    out = np.zeros(n, dtype=rp.PAIR_DTYPE)
    axis = rng.normal(size=(n, 3))
    axis /= np.linalg.norm(axis, axis=1)[:, None]
    ang = rng.uniform(-rot_step, rot_step, size=n)
    K = np.zeros((n, 3, 3))
    K[:, 0, 1], K[:, 0, 2], K[:, 1, 0] = -axis[:, 2], axis[:, 1], axis[:, 2]
    K[:, 1, 2], K[:, 2, 0], K[:, 2, 1] = -axis[:, 0], -axis[:, 1], axis[:, 0]
    rot = np.eye(3)[None] + np.sin(ang)[:, None, None] * K + (1 - np.cos(ang))[:, None, None] * (K @ K)
    out["pos1"] = shapes[i, :3] + rng.uniform(-trans_step, trans_step, size=(n, 3))
    out["orientation1"] = (rot @ shapes[i, 3:].reshape(n, 3, 3)).reshape(n, 9)
    out["idx1"] = 0
    out["pos2"] = shapes[j, :3]; out["orientation2"] = shapes[j, 3:]; out["idx2"] = 0
    out["translation"] = -np.round((shapes[j, :3] - shapes[i, :3]) / L) * L
    return out


def threaded_verdicts(traits, pairs, threads):
    parts = np.array_split(np.arange(pairs.shape[0]), threads)
    with ThreadPoolExecutor(threads) as ex:
        res = list(ex.map(lambda idx: traits.pair_test_batch(pairs[idx[0]:idx[-1] + 1])[0] if len(idx) else
                          np.zeros(0, np.uint8), parts))
    return np.concatenate(res)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--target", type=float, default=1e8)
    ap.add_argument("--workload", default="sc_1m")
    ap.add_argument("--equil", type=int, default=100)
    ap.add_argument("--ref-cycles", type=int, default=2)
    ap.add_argument("--chunk", type=int, default=4_000_000)
    args = ap.parse_args()
    threads = len(os.sched_getaffinity(0))
    desc = rp.make_desc(hard=rp.RP_HARD_SPHEROCYLINDER, radius=W.SC_RADIUS, length=W.SC_LENGTH)
    strict, native = R.RefTraits(desc), (R.RefTraits(desc, native=True) if R.have_ref(native=True) else None)
    assert R.have_ref(), "needs oracle/_ref/librampack_ref.so (the compiled reference)"
    dev = rp.Device(0, seed=2024)
    dev.set_interaction(desc)
    box, shapes = W.sc_config(args.workload)
    dev.upload(box, shapes)
    for c in range(args.equil):
        dev.sweep(rp.RP_MOVE_ROTOTRANSLATION, 0.2, 0.15, c)
    box, shapes = dev.download()
    rng_range = W.SC_LENGTH + 2 * W.SC_RADIUS
    contact2 = (2 * W.SC_RADIUS) ** 2
    total = mismatches = overlapping = fma_sensitive = 0
    near, snapshots = [], 0
    t_dev = t_ref = 0.0
    t0 = time.time()
    rng = np.random.default_rng(20240)
    while total < args.target:
        # a snapshot written by the reference's own Simulation (NVT: box_scaler 0)
        dom = (1, 1, min(16, threads))
        _, box, shapes = R.ref_run_simulation(native or strict, box, shapes, 1000 + snapshots, 1.0, 1.0, 0.2, 0.15, 0,
                                              0.0, dom, dom[2], args.ref_cycles, 0)
        snapshots += 1
        ij = neighbour_pairs(box, shapes, rng_range)
        print(f"snapshot {snapshots}: {ij.shape[0]} pairs within range ({time.time() - t0:.0f} s)", file=sys.stderr,
              flush=True)
        for c0 in range(0, ij.shape[0], args.chunk):
            if total >= args.target:
                break
            pairs = build_pairs(box, shapes, ij[c0:c0 + args.chunk], rng)
            t = time.time(); got, _ = dev.pair_test_batch(pairs); t_dev += time.time() - t
            t = time.time(); want = threaded_verdicts(strict, pairs, threads); t_ref += time.time() - t
            bad = np.nonzero(got != want)[0]
            mismatches += bad.size
            overlapping += int(want.sum())
            if native is not None:
                fma_sensitive += int((threaded_verdicts(native, pairs, threads) != want).sum())
            p1 = pairs["pos1"]; p2 = pairs["pos2"] + pairs["translation"]
            h1 = pairs["orientation1"][:, [2, 5, 8]] * (W.SC_LENGTH / 2)
            h2 = pairs["orientation2"][:, [2, 5, 8]] * (W.SC_LENGTH / 2)
            d2 = seg_dist2(p1, h1, p2, h2)
            for k in np.nonzero(np.abs(d2 - contact2) <= 1e-12 * contact2)[0]:
                near.append({"snapshot": snapshots, "i": int(ij[c0 + k, 0]), "j": int(ij[c0 + k, 1]),
                             "d2_minus_contact2": float(d2[k] - contact2), "device": int(got[k]),
                             "reference": int(want[k])})
            for k in bad[:20]:
                near.append({"MISMATCH": True, "snapshot": snapshots, "i": int(ij[c0 + k, 0]),
                             "j": int(ij[c0 + k, 1]), "d2": float(d2[k]), "device": int(got[k]),
                             "reference": int(want[k])})
            total += pairs.shape[0]
    out = {"pairs_tested": int(total), "snapshots": snapshots, "workload": args.workload,
           "verdict_mismatches": int(mismatches), "overlapping_pairs": int(overlapping),
           "pairs_within_1e-12_of_contact": near, "fma_sensitive_pairs_strict_vs_shipped_flags": int(fma_sensitive),
           "device_seconds": t_dev, "reference_seconds": t_ref, "reference_threads": threads,
           "checker": "oracle/_ref/librampack_ref.so (unmodified reference, -O3 -ffp-contract=off)",
           "wall_seconds": time.time() - t0}
    print(json.dumps(out))
    dev.close()
    return 0 if mismatches == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
