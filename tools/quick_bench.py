"""Development probe (not the contract bench): times rp_sweep / rp_try_scaling on the spherocylinder lattice."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import rampack_b200 as rp  # noqa: E402
import helpers as H  # noqa: E402


def main():
    sizes = {"12k": (28, 28, 16), "100k": (56, 56, 32), "1M": (120, 120, 70)}
    which = sys.argv[1:] or ["100k"]
    dev = rp.Device(0, seed=1234)
    for name in which:
        box, shapes = H.sc_lattice(sizes[name])
        n = shapes.shape[0]
        dev.set_interaction(H.desc_sc())
        t0 = time.time()
        dev.upload(box, shapes)
        print(f"[{name}] N={n} upload+grid {time.time() - t0:.3f}s grid={tuple(dev.grid_info().cell_divisions)}",
              flush=True)
        therm = int(os.environ.get("THERM", "30"))
        for cycle in range(therm):
            c = dev.sweep(rp.RP_MOVE_ROTOTRANSLATION, 0.05, 0.05, cycle)
        ms = []
        acc = 0
        for cycle in range(therm, therm + 10):
            c = dev.sweep(rp.RP_MOVE_ROTOTRANSLATION, 0.05, 0.05, cycle)
            k = dev.last_kernel_ms()
            ms.append((k["sweep"], k["sweep_grid"]))
            acc += c.accepted
        sw = np.median([m[0] for m in ms])
        gr = np.median([m[1] for m in ms])
        print(f"[{name}] sweep kernels {sw:.3f} ms, sweep grid {gr:.3f} ms -> {n / (sw + gr) * 1e3:.3e} moves/s "
              f"(kernels only {n / sw * 1e3:.3e}); acceptance {acc / (10 * n):.3f}; "
              f"rej_domain {c.rejected_domain / n:.4f}", flush=True)
        t0 = time.time()
        ov = dev.count_total_overlaps(False)
        print(f"[{name}] count_total_overlaps={ov} in {time.time() - t0:.4f}s (incl. grid rebuild "
              f"{dev.last_kernel_ms()['rebuild']:.3f} ms)", flush=True)
        t0 = time.time()
        ov = dev.count_total_overlaps(True)
        print(f"[{name}] count_total_overlaps(early)={ov} in {time.time() - t0:.4f}s", flush=True)
        dE = dev.try_scaling(box * 0.999)
        print(f"[{name}] try_scaling(0.999) dE={dE} {dev.last_kernel_ms()['scaling']:.3f} ms", flush=True)
        dev.revert_scaling()
        dE = dev.try_scaling(box * 1.001)
        print(f"[{name}] try_scaling(1.001) dE={dE} {dev.last_kernel_ms()['scaling']:.3f} ms", flush=True)
        dev.revert_scaling()


if __name__ == "__main__":
    main()
