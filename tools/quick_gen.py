"""Development probe for the generic sweep path: times rp_sweep on one of the EXTRA_CONFIGS of rampack_b200/workloads.py."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rampack_b200 as rp  # noqa: E402
from rampack_b200 import workloads as W  # noqa: E402


def main():
    which = sys.argv[1:] or ["kmer4_200k"]
    therm = int(os.environ.get("THERM", "5"))
    reps = int(os.environ.get("REPS", "5"))
    counting = bool(int(os.environ.get("COUNTING", "0")))
    for name in which:
        cfg = W.EXTRA_CONFIGS[name]
        dev = rp.Device(0, seed=1234)
        box, shapes = W.extra_lattice(cfg if counting else dict(cfg, relax=None))
        n = shapes.shape[0]
        dev.set_interaction(cfg["desc"]())
        dev.upload(box, shapes)
        if counting:
            dev.toggle_overlap_counting(True)
        for cycle in range(therm):
            dev.sweep(rp.RP_MOVE_ROTOTRANSLATION, cfg["moves"][0], cfg["moves"][1], cycle)
        ms, acc = [], 0
        for cycle in range(therm, therm + reps):
            c = dev.sweep(rp.RP_MOVE_ROTOTRANSLATION, cfg["moves"][0], cfg["moves"][1], cycle)
            k = dev.last_kernel_ms()
            ms.append((k["sweep"], k["sweep_grid"]))
            acc += c.accepted
        sw = np.median([m[0] for m in ms])
        gr = np.median([m[1] for m in ms])
        print(f"[{name}] N={n} counting={counting} sweep kernels {sw:.3f} ms, sweep grid {gr:.3f} ms -> "
              f"{n / (sw + gr) * 1e3:.3e} moves/s; acceptance {acc / (reps * n):.3f}", flush=True)
        dev.close()


if __name__ == "__main__":
    main()
