"""Generates tests/golden/sc_l5_nematic_start.npz: an EQUILIBRATED nematic configuration of hard spherocylinders
L/D = 5 at p = 1.3 (N = 363) that both sides of the statistical parity test start from.  A lattice start does not do:
the density relaxes over > 10^5 cycles (one box move per cycle), far more than a test can afford, and the average over
a drifting window measures the relaxation rate of the sampler, not the equilibrium.  Produced with the plain-C
restatement's sequential Monte Carlo loop (oracle/rp_oracle.c, rpo_run_cycles): run HERE,
    python tests/golden/make_nematic_start.py
The printed block densities show where the drift ends."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import stat_cases as S  # noqa: E402
from oracle import refwrap as R  # noqa: E402


def main():
    seed = int(sys.argv[1]) if len(sys.argv) > 1 else 7
    blocks = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    case = S.CASES["sc_l5_dense"]
    box, shapes = S.sc_columns((11, 11, 3), (1.2, 1.2, 6.5))
    pk = R.PortPacking(R.PortTraits(case["desc"]()), box, shapes)
    for b in range(blocks):
        dens = []
        for k in range(10):
            pk.run_cycles(seed * 100000 + b * 10 + k, 1000, case["trans_step"], case["rot_step"], 1.0,
                          case["pressure"], True, 0.012)
            bx, sh = pk.get()
            dens.append(S.number_density(bx, sh))
        print(f"block {b} ({(b + 1) * 10000} cycles): density {np.mean(dens):.5f}  P2 {S.nematic_p2(sh):.3f}  "
              f"tau {S.smectic_tau(bx, sh):.3f}", flush=True)
        np.savez(os.path.join(HERE, f"sc_l5_nematic_start_seed{seed}.npz"), box=bx, shapes=sh, cycles=(b + 1) * 10000)


if __name__ == "__main__":
    main()
