"""Equilibrates the ORDERED L/D = 5 spherocylinder state points of the statistical-parity test on the device and writes
them as starting snapshots (tests/golden/sc_l5_ordered.npz): a nematic and a smectic point of BASELINE config 2.

From a lattice start neither side reaches these states within a test's budget (tests/stat_cases.py), so the state is
equilibrated ONCE here (run on the GPU box: `python tests/golden/make_ordered_snapshots.py`, output under gpurun_out/,
then moved to tests/golden/) and BOTH sides — the reference run of make_stat_golden.py and the device run of
test_gpu_statistics.py — start from the same committed snapshot with the step sizes adapted here.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import rampack_b200 as rp  # noqa: E402
import stat_cases as S  # noqa: E402

POINTS = {"sc_l5_nematic": 1.4, "sc_l5_smectic": 2.2}
CELLS, CELL = (14, 14, 4), (1.2, 1.2, 6.5)


def main():
    cycles = int(os.environ.get("EQ_CYCLES", "400000"))
    out, log = {}, {}
    only = os.environ.get("EQ_POINTS")
    for name, pressure in POINTS.items():
        if only and name not in only.split(","):
            continue
        pressure = float(os.environ.get("EQ_PRESSURE_" + name, pressure))
        dev = rp.Device(0, seed=20261105)
        dev.set_interaction(rp.make_desc(hard=rp.RP_HARD_SPHEROCYLINDER, radius=0.5, length=5.0))
        box, shapes = S.sc_columns(CELLS, CELL) if "nematic" in name else S.sc_lattice(CELLS, CELL)
        dev.upload(box, shapes)
        sim = rp.NativeSimulation(dev, [(rp.RP_MOVE_ROTOTRANSLATION, 0.1, 0.1)], 1.0, pressure,
                                  ("delta_triclinic", 0.01, False), seed=777)
        trace = []
        for block in range(20):
            sim.run(cycles // 20, adjust_step_sizes=True)
            b, s = dev.download()
            trace.append({"density": S.number_density(b, s), "p2": S.nematic_p2(s), "tau": S.smectic_tau(b, s)})
            print(name, block, trace[-1], flush=True)
        st = sim.stats()
        b, s = dev.download()
        assert dev.count_total_overlaps(False) == 0
        out[name + "_box"] = b
        out[name + "_shapes"] = s
        log[name] = {"pressure": pressure, "cycles": cycles, "trans_step": st.translation_step[0],
                     "rot_step": st.rotation_step[0], "box_step": st.scaler_step, "trace": trace}
        dev.close()
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    tag = os.environ.get("EQ_TAG", "")
    np.savez_compressed(os.path.join(ROOT, "gpurun_out", f"sc_l5_ordered{tag}.npz"), **out)
    json.dump(log, open(os.path.join(ROOT, "gpurun_out", f"sc_l5_ordered{tag}.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
