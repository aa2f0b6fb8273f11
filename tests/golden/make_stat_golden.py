"""Generates tests/golden/stat_golden.json: the statistical-parity cases of tests/stat_cases.py run with the
UNMODIFIED reference (oracle/_ref/librampack_ref_native.so = Simulation::integrate of /root/reference, shipped
optimisation level).  Run HERE:  python tests/golden/make_stat_golden.py [case ...]

Protocol per case: `therm` cycles, then `avg` cycles in blocks of `every` cycles; after each block the configuration is
read back and the observables of stat_cases.observables() are evaluated.  Every block is one Simulation::integrate
call continuing from the previous block's final packing (fresh seed per block, fixed step sizes — thermalisation-phase
step adaptation is off because each call has thermCycles = 0)."""
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import stat_cases as S  # noqa: E402
from oracle import refwrap as R  # noqa: E402

OUT = os.path.join(HERE, "stat_golden.json")
SCALER = {"delta_v": 1, "delta_triclinic": 2}


def run_case(name, seed=4321):
    case = S.CASES[name]
    traits = R.RefTraits(case["desc"](), native=True)
    box, shapes = case["state"]()
    kind, step = case["scaler"][:2]
    assert kind != "delta_triclinic" or case["scaler"][2] is False     # the harness runs scaleTogether = true
    args = (case["temperature"], case["pressure"], case["trans_step"], case["rot_step"], SCALER[kind], step,
            (1, 1, 1), 1)
    t0 = time.time()
    res, box, shapes = R.ref_run_simulation(traits, box, shapes, seed, *args, 0, case["therm"], case["therm"])
    series = {}
    acc_m = acc_s = att_m = att_s = 0
    for b in range(case["avg"] // case["every"]):
        res, box, shapes = R.ref_run_simulation(traits, box, shapes, seed + 1 + b, *args, 0, case["every"],
                                                case["every"])
        acc_m += res.moves_accepted; att_m += res.moves_attempted
        acc_s += res.scalings_accepted; att_s += res.scalings_attempted
        for k, v in S.observables(case, box, shapes).items():
            series.setdefault(k, []).append(v)
    out = {"seconds": time.time() - t0, "samples": len(series["density"]),
           "move_acceptance": acc_m / max(1, att_m), "scaling_acceptance": acc_s / max(1, att_s)}
    for k, v in series.items():
        mean, se = S.block_stats(v)
        out[k] = {"mean": mean, "se": se}
    return out


REF_SEEDS = [4321, 55, 77, 99, 1234, 31337, 2718, 16180]


def run_case_all_seeds(name):
    """Independent reference runs (in parallel processes) combined with stat_cases.combine_runs."""
    from concurrent.futures import ProcessPoolExecutor
    k = S.CASES[name].get("seeds_ref", 1)
    with ProcessPoolExecutor(k) as ex:
        runs = list(ex.map(run_case, [name] * k, REF_SEEDS[:k]))
    keys = [key for key, v in runs[0].items() if isinstance(v, dict)]
    comb = S.combine_runs([{key: (r[key]["mean"], r[key]["se"]) for key in keys} for r in runs])
    out = {"runs": runs, "seeds": REF_SEEDS[:k], "samples": runs[0]["samples"],
           "move_acceptance": runs[0]["move_acceptance"], "scaling_acceptance": runs[0]["scaling_acceptance"]}
    for key, (mean, se) in comb.items():
        out[key] = {"mean": mean, "se": se}
    return out


def main():
    names = sys.argv[1:] or list(S.CASES)
    golden = json.load(open(OUT)) if os.path.exists(OUT) else {}
    for name in names:
        golden[name] = run_case_all_seeds(name)
        print(name, json.dumps(golden[name]), flush=True)
        json.dump(golden, open(OUT, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
