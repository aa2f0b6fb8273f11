#!/usr/bin/env python
"""bench.py — MC particle moves/s (and volume moves/s) of the device-resident NpT sweep, hard spherocylinders L/D = 5.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload sc_1m]

A "step" is one NpT cycle over the whole packing: one bulk trial-move pass (N rototranslation attempts, rp_sweep) plus
one box-move proposal (Simulation::tryScaling semantics, rp_try_scaling / rp_revert_scaling).
  value  = attempted particle moves / device time of K steps, state resident in HBM        (CUDA events, max over ranks)
  e2e    = the same through the C ABI with HOST buffers: rp_upload (pinned host -> device) + cycle + rp_download
N > 1 ranks (--parallelism):
  sharded  (default) ONE packing of the named workload, the sweep spatially sharded over the ranks (z-slabs of cell
           layers; k_sweep writes accept bytes and boundary-layer records into the peers' memory over NVLink, flag
           kernels order the colours; box-move recount split with a device all-reduce) — bit-identical to the 1-GPU
           run; "scaling": "strong"
  replicas independent packings (the pressure points of a scan, north_star), one per GPU, no data-path collective;
           "scaling": "weak"
--impl reference times the reference's own OpenMP CPU implementation (oracle/_ref, unmodified sources) on the host.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "mc_particle_moves_per_s_hard_sc"
UNIT = "moves/s"
TRANS_STEP, ROT_STEP = 0.05, 0.05      # rototranslation(trans_step=0.05, rot_step=0.05) before adaptation
TEMPERATURE, PRESSURE = 1.0, 1.2
SEED = 1234
THERM_DEFAULT = 2000                   # thermalisation cycles with step-size adaptation (Simulation.cpp:128-135)
# flops of one pair test in the reference arithmetic: centre-distance pre-filter, full segment test
F_PRE, F_FULL = 14, 105
B_MOVE = 200                           # algorithmic bytes per attempted move: state read + write-back + list entry
CULL_INSTR_PER_PAIR = 10.0             # issue slots of the FP32 capsule cull per (trial, candidate) pair and lane (SASS)


def box_step_for(n):
    """Start value of the delta_triclinic step.  A uniform compression by eps brings every near-contact pair closer by
    eps * r: with ~10 N pairs per unit gap an overlap-free proposal needs eps ~ 1 / (10 N), i.e. a step of a few 1 / N
    for these boxes; the thermalisation adapts it from there (Simulation.cpp:598-607)."""
    return min(0.01, 3.0 / n)


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.path = f"/tmp/rp_clocks_{os.getpid()}.csv"

    def start(self):
        try:
            self.out = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=self.out,
                                         stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.out.close()
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        try:
            os.remove(self.path)
        except OSError:
            pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(smax), "power_w_max": max(power),
                "samples": len(sm), "reasons": sorted(reasons)}


def choose_domains(cores, box_heights, total_range, cell):
    """Largest a*b*c <= cores with every domain wider than ghost layer + one grid cell
    (DomainDecomposition.cpp:54-61: domain - ghost > ngCell, ghost = totalRange - range + ngCell ~ one cell)."""
    limit = [max(1, int(h / (2.2 * cell + 1e-9))) for h in box_heights]
    best, best_p = (1, 1, 1), 1
    for a in range(1, min(limit[0], cores) + 1):
        for b in range(1, min(limit[1], cores // a) + 1):
            c = min(limit[2], cores // (a * b))
            if c >= 1 and a * b * c > best_p:
                best, best_p = (a, b, c), a * b * c
    return best


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference_cycles(box, shapes, cycles, seed, threads=None, steps=None, desc=None, pressure=PRESSURE,
                         scaler=2, total_range=6.0):
    """`cycles` NpT cycles of the reference's own Simulation (oracle/_ref when built, else the single-thread port).
    steps = (translation step, rotation step, box step) — frozen: the averaging phase does not adapt them.
    Returns (moves attempted, scalings attempted, seconds for moves + scaling + domain decomposition, info, state)."""
    from oracle import refwrap as R
    from rampack_b200 import workloads as W
    import rampack_b200 as rp
    if desc is None:
        desc = rp.make_desc(hard=rp.RP_HARD_SPHEROCYLINDER, radius=W.SC_RADIUS, length=W.SC_LENGTH)
    n = shapes.shape[0]
    t_step, r_step, b_step = steps if steps is not None else (TRANS_STEP, ROT_STEP, box_step_for(n))
    if R.have_ref(native=True) or R.have_ref():
        native = R.have_ref(native=True)
        traits = R.RefTraits(desc, native=native)
        cores = threads or host_cores()
        heights = [box[0, 0], box[1, 1], box[2, 2]]
        dom = choose_domains(cores, heights, total_range, total_range * 1.04)
        nthreads = dom[0] * dom[1] * dom[2]
        # the whole run is the "averaging" phase: the reference adapts step sizes in the thermalisation phase only
        res, box_out, shapes_out = R.ref_run_simulation(traits, box, shapes, seed, TEMPERATURE, pressure, t_step,
                                                        r_step, scaler, b_step, dom, nthreads, 0, cycles)
        seconds = res.move_seconds + res.scaling_seconds + res.domain_decomposition_seconds
        info = {"kind": "reference", "cores": nthreads, "domains": list(dom),
                "flags": "-O3 -march=x86-64-v3 -fopenmp" if native else "-O3 -ffp-contract=off -fopenmp",
                "move_seconds": res.move_seconds, "scaling_seconds": res.scaling_seconds,
                "scalings_attempted": int(res.scalings_attempted)}
        return int(res.moves_attempted), int(res.scalings_attempted), seconds, info, (box_out, shapes_out)
    pk = R.PortPacking(R.PortTraits(desc), box, shapes)
    res = pk.run_cycles(seed, cycles, t_step, r_step, TEMPERATURE, pressure, True, b_step)
    info = {"kind": "port", "cores": 1, "domains": [1, 1, 1], "flags": "-O2 -ffp-contract=off",
            "move_seconds": res.move_seconds, "scaling_seconds": res.scaling_seconds,
            "scalings_attempted": int(res.scalings_attempted)}
    b, s = pk.get()
    return int(res.moves_attempted), int(res.scalings_attempted), res.move_seconds + res.scaling_seconds, info, (b, s)


def workload_string(name, n, therm, steps3, sharded=False, per_gpu=False):
    """The one description both arms print (`config.workload`): same lattice, same thermalisation, same frozen steps."""
    where = " in ONE packing sharded over all GPUs" if sharded else (" per GPU" if per_gpu else "")
    return (f"{name}: hard spherocylinders l=5 r=0.5, N={n}{where}, sc lattice eta=0.475 + {therm} thermalisation "
            f"cycles on the device with the reference's step adaptation, then frozen steps: NpT T=1 p={PRESSURE}, "
            f"rototranslation({steps3[0]:.6g}, {steps3[1]:.6g}), delta_triclinic({steps3[2]:.6g})")


def thermalised_state(workload, therm, local=0, seed=SEED, pressure=PRESSURE, stream=None, comm=None):
    """Lattice -> `therm` NpT cycles on the device with step-size adaptation (Simulation.cpp:128-135, 533-607).
    Deterministic for a given (workload, therm, seed): both arms call it and start from the same state with the same
    frozen step sizes.  Returns (device, simulation, (translation, rotation, box) steps)."""
    import rampack_b200 as rp
    from rampack_b200 import workloads as W
    from rampack_b200 import NativeSimulation
    dev = rp.Device(device=local, seed=seed, cuda_stream=stream)
    if comm is not None:
        dev.comm_init(*comm)
    desc = rp.make_desc(hard=rp.RP_HARD_SPHEROCYLINDER, radius=W.SC_RADIUS, length=W.SC_LENGTH)
    box, shapes = W.sc_config(workload)
    dev.set_interaction(desc)
    dev.upload(box, shapes)
    sim = NativeSimulation(dev, [(rp.RP_MOVE_ROTOTRANSLATION, TRANS_STEP, ROT_STEP)], TEMPERATURE, pressure,
                           ("delta_triclinic", box_step_for(shapes.shape[0]), False), seed=seed)
    sim.run(therm, adjust_step_sizes=True)
    st = sim.stats()
    return dev, sim, (st.translation_step[0], st.rotation_step[0], st.scaler_step)


def main_reference(args):
    """The reference's own OpenMP Simulation on the host cores, from the state and with the step sizes the GPU arm is
    timed on (prepared, untimed, by the same deterministic device thermalisation; without a GPU: from the lattice)."""
    rank = env_int("RANK", 0)
    if rank != 0:
        return 0
    from rampack_b200 import workloads as W
    name = args.workload
    box, shapes = W.sc_config(name)
    est = shapes.shape[0] / (3.0e4 * max(1, min(host_cores(), 512))) * (args.steps + args.warmup)
    if est > 150 and name != "sc_100k":        # K + W cycles of the metric's configuration must fit a few minutes
        name = "sc_100k"
    n = W.sc_config(name)[1].shape[0]
    steps3, same_state = (TRANS_STEP, ROT_STEP, box_step_for(n)), False
    box, shapes = W.sc_config(name)
    try:
        import torch
        if torch.cuda.is_available():
            dev, sim, steps3 = thermalised_state(name, args.therm)
            box, shapes = dev.download()
            sim.close()
            dev.close()
            same_state = True
    except Exception as ex:        # no device / no library: the reference arm still runs, from the lattice
        print(f"bench.py --impl reference: device thermalisation unavailable ({ex}); starting from the lattice",
              file=sys.stderr)
    state = (box, shapes)
    if args.warmup > 0:
        _, _, _, _, state = run_reference_cycles(box, shapes, args.warmup, SEED, steps=steps3)
    moves, scalings, seconds, info, _ = run_reference_cycles(state[0], state[1], args.steps, SEED + 1, steps=steps3)
    value = moves / seconds
    sample = f"{args.steps} NpT cycles of {name} (N={n}) after {args.warmup} warm-up cycles, all particles"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": seconds / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_string(name, n, args.therm if same_state else 0, steps3),
                       "parallelism": f"OpenMP domains {info['domains']}", "same_state_as_gpu_arm": same_state},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": info["cores"], "kind": info["kind"],
                             "sample": sample, "flags": info["flags"]},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "volume_moves_per_s": (scalings / info["scaling_seconds"]) if info["scaling_seconds"] > 0 else None,
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


def sweep_cells(box, n, total_range):
    """Cells of the checkerboard grid rp_sweep builds (rampack_b200/csrc/rp_sweep.cu): width >= totalRange and wide
    enough for ~16 particles, even counts."""
    heights = [np.linalg.det(box) / np.linalg.norm(np.cross(box[:, (i + 1) % 3], box[:, (i + 2) % 3]))
               for i in range(3)]
    min_width = total_range * (1 + 1e-9)
    width = max(min_width, (16.0 * abs(np.linalg.det(box)) / n) ** (1.0 / 3.0))
    cells = []
    for h in heights:
        c = max(2, int(h / width))
        if c & 1:           # small grids round up when the narrower cell is admissible (rp_sweep.cu)
            c = c + 1 if c <= 7 and h / (c + 1) >= min_width else c - 1
        cells.append(c)
    return cells


def measure_extra(name, cfg, local, stream, peak_gflops, steps, budget_s=6.0, cpu=True):
    """One of the other BASELINE configurations (rampack_b200/workloads.py EXTRA_CONFIGS): thermalise with step
    adaptation, time `steps` cycles with CUDA events, and put the reference's CPU figure for a bounded sample of the
    same lattice (same density, same step sizes) beside it."""
    import torch
    import rampack_b200 as rp
    from rampack_b200 import workloads as W
    from rampack_b200 import NativeSimulation
    out = {}
    desc = cfg["desc"]()
    box, shapes = W.extra_lattice(cfg)
    n = shapes.shape[0]
    dev = rp.Device(device=local, seed=SEED, cuda_stream=stream.cuda_stream)
    dev.set_interaction(desc)
    dev.upload(box, shapes)
    sim = NativeSimulation(dev, [(rp.RP_MOVE_ROTOTRANSLATION,) + tuple(cfg["moves"])], TEMPERATURE, cfg["p"],
                           cfg["scaler"], seed=SEED)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed(cycles, adjust=False):
        s0 = sim.stats()
        with torch.cuda.stream(stream):
            e0.record(stream)
            sim.run(cycles, adjust_step_sizes=adjust)
            e1.record(stream)
        torch.cuda.synchronize()
        s1 = sim.stats()
        return e0.elapsed_time(e1), s0, s1

    if cfg.get("relax"):
        # overlap_relaxation run (Simulation::relaxOverlaps, Simulation.cpp:218-323): overlap counting on (with the
        # helper interaction when the configuration names one), cycles until the cached count reaches zero
        helper = cfg.get("relax_desc")
        if helper:
            sim.close()
            dev.set_interaction(helper())
            dev.upload(box, shapes)
            sim = NativeSimulation(dev, [(rp.RP_MOVE_ROTOTRANSLATION,) + tuple(cfg["moves"])], TEMPERATURE, cfg["p"],
                                   cfg["scaler"], seed=SEED)
        dev.toggle_overlap_counting(True)
        start = dev.cached_overlaps()
        ms_total, cycles, moves = 0.0, 0, 0
        while dev.cached_overlaps() > 0 and ms_total < 3 * budget_s * 1e3 and cycles < 5000:
            ms, s0, s1 = timed(5, adjust=True)
            ms_total += ms
            cycles += 5
            moves += s1.moves_attempted[0] - s0.moves_attempted[0]
        left = dev.cached_overlaps()
        out["overlap_relaxation"] = {"overlaps_start": int(start), "overlaps_end": int(left), "cycles": cycles,
                                     "value": moves / (ms_total * 1e-3) if ms_total > 0 else None, "unit": UNIT,
                                     "ms_per_cycle": ms_total / max(1, cycles), "helper": bool(helper),
                                     "finished": left == 0}
        if left == 0:
            # the production run continues with the plain hard shape from the relaxed packing
            dev.toggle_overlap_counting(False)
            if helper:
                st = sim.stats()
                rbox, rshapes = dev.download()
                sim.close()
                dev.set_interaction(desc)
                dev.upload(rbox, rshapes)
                sim = NativeSimulation(dev, [(rp.RP_MOVE_ROTOTRANSLATION, st.translation_step[0],
                                              st.rotation_step[0])], TEMPERATURE, cfg["p"],
                                       (cfg["scaler"][0], st.scaler_step) + tuple(cfg["scaler"][2:]), seed=SEED + 1)
        else:
            # not relaxed within the time budget (the last few overlaps of 2e5 particles take thousands of cycles): the
            # production cycles below are timed on the UNCOMPRESSED lattice of the same configuration instead, with the
            # plain hard shape and overlap counting off, so that they measure the ordinary sweep and the reference's
            # integration run accepts the same start
            out["overlap_relaxation"]["finished"] = False
            out["production_start"] = "uncompressed lattice (the relaxation did not finish within the time budget)"
            sim.close()
            dev.toggle_overlap_counting(False)
            dev.set_interaction(desc)
            box, shapes = W.extra_lattice(dict(cfg, relax=None))
            dev.upload(box, shapes)
            sim = NativeSimulation(dev, [(rp.RP_MOVE_ROTOTRANSLATION,) + tuple(cfg["moves"])], TEMPERATURE, cfg["p"],
                                   cfg["scaler"], seed=SEED + 2)
    # thermalisation with the reference's step adaptation, bounded by the time budget
    timed(2, adjust=True)             # (first calls: allocations)
    ms, _, _ = timed(3, adjust=True)
    therm = int(max(10, min(cfg.get("therm", 200), budget_s * 1e3 / max(ms / 3, 1e-3))))
    timed(therm, adjust=True)
    timed(3)
    launches0 = dev.launch_count()
    ms, s0, s1 = timed(steps)
    attempted = s1.moves_attempted[0] - s0.moves_attempted[0]
    accepted = s1.moves_accepted[0] - s0.moves_accepted[0]
    evaluated = s1.scalings_evaluated - s0.scalings_evaluated
    scaling_ms = s1.scaling_ms - s0.scaling_ms
    sweep_ms = (s1.sweep_kernel_ms - s0.sweep_kernel_ms) + (s1.sweep_grid_ms - s0.sweep_grid_ms)
    steps3 = (s1.translation_step[0], s1.rotation_step[0], s1.scaler_step)
    k_cand, k_full = dev.neighbour_stats()
    n_centres = max(1, desc.n_centres)
    f_pre, f_full = cfg["flops"]
    w_move = n_centres * (k_cand * f_pre + k_full * f_full)
    value = attempted / (ms * 1e-3)
    achieved = attempted * w_move / (max(sweep_ms, 1e-9) * 1e-3) * 1e-12
    box_now = np.array(list(s1.box)).reshape(3, 3)
    out.update({"value": value, "unit": UNIT, "n": n, "interaction_centres": n * n_centres, "steps": steps,
                "thermalisation_cycles": therm, "ms_per_step": ms / steps,
                "acceptance": accepted / max(1, attempted),
                "step_sizes": {"translation": steps3[0], "rotation": steps3[1], "box": steps3[2]},
                "number_density": n / abs(np.linalg.det(box_now)),
                "box_moves_per_s": (evaluated / (scaling_ms * 1e-3)) if scaling_ms > 0 and evaluated else None,
                "box_moves": {"attempted": steps, "evaluated": int(evaluated),
                              "accepted": int(s1.scalings_accepted - s0.scalings_accepted)},
                "gpu_launches": int(dev.launch_count() - launches0),
                "roofline": {"bound": "fp64", "achieved": achieved, "peak": peak_gflops * 1e-3, "unit": "TFLOP/s",
                             "frac": achieved / (peak_gflops * 1e-3), "flop_per_move": w_move, "k_cand": k_cand,
                             "k_within_range": k_full, "flops_per_pair": [f_pre, f_full],
                             "sweep_ms_per_step": sweep_ms / steps,
                             "hbm_frac": attempted * B_MOVE / (max(sweep_ms, 1e-9) * 1e-3) * 1e-9 / 6540.0}})
    if "nematic" in cfg.get("observe", ()):
        out["nematic_p2"] = float(dev.nematic_order()[1])
    if cpu:
        try:
            from oracle import refwrap as R
            cbox, cshapes = W.extra_lattice(dict(cfg, relax=None), cells=cfg["cpu_cells"])
            # the CPU sample runs at the density the device run has reached (same per-move work)
            scale = (n / abs(np.linalg.det(box_now)) / (cshapes.shape[0] / abs(np.linalg.det(cbox)))) ** (-1.0 / 3.0)
            cbox, cshapes = cbox * scale, cshapes.copy()
            cshapes[:, :3] *= scale
            native = R.have_ref(native=True)
            tr = R.RefTraits(desc, native=native)
            total_range = 2 * tr.total_range_radius
            scaler_kind = 1 if cfg["scaler"][0] == "delta_v" else 2
            t0 = time.perf_counter()
            m1, _, sec1, info, state = run_reference_cycles(cbox, cshapes, 1, SEED, steps=steps3, desc=desc,
                                                            pressure=cfg["p"], scaler=scaler_kind,
                                                            total_range=total_range)
            cycles = int(max(1, min(50, (budget_s - (time.perf_counter() - t0)) / max(sec1, 1e-4))))
            m, sc, sec, info, _ = run_reference_cycles(state[0], state[1], cycles, SEED + 1, steps=steps3, desc=desc,
                                                       pressure=cfg["p"], scaler=scaler_kind, total_range=total_range)
            out["cpu_baseline"] = {"value": m / sec, "unit": UNIT, "cores": info["cores"], "kind": info["kind"],
                                   "sample": f"{cycles} NpT cycles of the same lattice at N={cshapes.shape[0]} "
                                             f"(same density and step sizes), domains {info['domains']}",
                                   "box_moves_per_s": (sc / info["scaling_seconds"]) if info["scaling_seconds"] > 0
                                   else None}
        except Exception as ex:
            out["cpu_baseline"] = {"value": None, "kind": "unavailable", "sample": str(ex)}
    sim.close()
    dev.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="sc_1m", choices=["sc_12k", "sc_100k", "sc_1m", "sc_4m"])
    ap.add_argument("--therm", type=int, default=THERM_DEFAULT,
                    help="untimed thermalisation cycles (with step-size adaptation) before the warm-up")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-cycles", type=int, default=1)
    ap.add_argument("--configs", default="all",
                    help="other BASELINE configurations measured after the metric's (N = 1 only): all, none, or a "
                         "comma-separated list of names from rampack_b200/workloads.py EXTRA_CONFIGS")
    ap.add_argument("--parallelism", default="sharded", choices=["sharded", "replicas"],
                    help="what N > 1 ranks do (ignored at N = 1)")
    args = ap.parse_args()
    if args.impl == "reference":
        return main_reference(args)

    import torch
    import rampack_b200 as rp
    from rampack_b200 import workloads as W

    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    sharded = world > 1 and args.parallelism == "sharded"
    rseed = SEED if sharded else SEED + rank          # sharded ranks are one simulation: same Philox key everywhere
    stream = torch.cuda.Stream()
    comm = None
    if sharded:
        ident = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            ident.copy_(torch.frombuffer(bytearray(rp.Device.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(ident, 0)
        comm = (bytes(ident.cpu().numpy().tobytes()), rank, world)
    # each replica is one pressure point of a scan (north_star): rank r runs p = 1.2 + 0.1 r.  The cycle loop is the
    # reference's Simulation::performCycle in C++ inside the library (rp_sim, rampack_b200/csrc/rp_sim.cu); the
    # thermalisation adapts the step sizes like the reference's, the timed cycles run with them frozen
    dev, sim, steps3 = thermalised_state(args.workload, args.therm, local, rseed,
                                         PRESSURE if sharded or world == 1 else PRESSURE + 0.1 * rank,
                                         stream.cuda_stream, comm)
    therm_box, therm_shapes = dev.download()
    n = therm_shapes.shape[0]

    # ---- device-resident timing: W warm-up + exactly K timed steps ----
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    sim.run(args.warmup)
    launches0 = dev.launch_count()
    s0 = sim.stats()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    with torch.cuda.stream(stream):
        e0.record(stream)
        sim.run(args.steps)
        e1.record(stream)
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    launches = dev.launch_count() - launches0
    s1 = sim.stats()
    # the timed region is a few tens of milliseconds: the clock sampler (100 ms period) keeps running over the warm-up
    # before it and the same cycles repeated after it
    if ms < 600:
        sim.run(max(1, int((600 - ms) / max(ms / args.steps, 1e-3))))
        torch.cuda.synchronize()
    clocks = sampler.stop() if rank == 0 else None
    sweep_ms = s1.sweep_kernel_ms - s0.sweep_kernel_ms
    grid_ms = s1.sweep_grid_ms - s0.sweep_grid_ms
    scaling_ms = s1.scaling_ms - s0.scaling_ms
    evaluated = s1.scalings_evaluated - s0.scalings_evaluated
    attempted = s1.moves_attempted[0] - s0.moves_attempted[0]
    accepted = s1.moves_accepted[0] - s0.moves_accepted[0]
    scalings_accepted = s1.scalings_accepted - s0.scalings_accepted
    # sharded: every rank reports the counters of the ONE packing; replicas: one packing per rank
    total_attempted = float(attempted) if sharded else sum_over_ranks(float(attempted))
    value = total_attempted / (ms * 1e-3)

    # ---- end to end through the C ABI with host buffers (pinned): upload + cycle + download every step ----
    # sharded: every rank moves its own slice of the packing over PCIe (rp_upload_slice all-gathers over NVLink,
    # rp_download_range reads the slice back); replicas / 1 GPU: the whole packing of the rank
    b_now, s_now = dev.download()
    first, count = rp.Device.slice_of(n, world, rank) if sharded else (0, n)
    host = torch.empty((count, 12), dtype=torch.float64).pin_memory().numpy()
    host[:] = s_now[first:first + count]
    e2e_box = b_now
    e2e_steps = max(3, min(args.steps, 10))
    e2e_pipeline = 1
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        if sharded:
            dev.upload_slice(e2e_box, host, first, n)
            sim.run(1)
            e2e_box = np.array(list(sim.stats().box)).reshape(3, 3)
            dev.download_range(first, count, out=host)
        else:
            dev.upload(e2e_box, host)
            sim.run(1)
            e2e_box, _ = dev.download(out=host)
    barrier()
    e2e_serial_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
    e2e_ms = e2e_serial_ms
    if not sharded:
        # The same steps with copies and compute overlapped (separate streams, separate copy engines): E2E_PIPELINE
        # packings in flight per GPU — e.g. the pressure points of a scan fed from the host — each on its own handle
        # and stream, each step still upload + cycle + download of its packing; a host thread per packing (the
        # library calls release the GIL).  H2D of one packing runs under the cycle of the second and the D2H of the
        # third.  The serial chain (one packing, nothing overlapped) stays in the line as e2e.serial_value.
        from concurrent.futures import ThreadPoolExecutor
        from rampack_b200 import NativeSimulation
        e2e_pipeline = max(1, env_int("E2E_PIPELINE", 4))
        st_now = sim.stats()
        lanes = [(dev, sim, host)]
        for k in range(1, e2e_pipeline):
            d2 = rp.Device(device=local, seed=rseed + 100 * k)
            d2.set_interaction(rp.make_desc(hard=rp.RP_HARD_SPHEROCYLINDER, radius=W.SC_RADIUS, length=W.SC_LENGTH))
            d2.upload(b_now, s_now)
            s2 = NativeSimulation(d2, [(rp.RP_MOVE_ROTOTRANSLATION, st_now.translation_step[0],
                                        st_now.rotation_step[0])], TEMPERATURE,
                                  PRESSURE if world == 1 else PRESSURE + 0.1 * rank,
                                  ("delta_triclinic", st_now.scaler_step, False), seed=rseed + 100 * k)
            h2 = torch.empty((n, 12), dtype=torch.float64).pin_memory().numpy()
            h2[:] = s_now
            s2.run(1)               # allocations, staging capacities
            lanes.append((d2, s2, h2))

        def lane_loop(lane):
            d, sm, hbuf = lane
            bx = b_now
            for _ in range(e2e_steps):
                d.upload(bx, hbuf)
                sm.run(1)
                bx, _ = d.download(out=hbuf)

        barrier()
        t0 = time.perf_counter()
        with ThreadPoolExecutor(max_workers=e2e_pipeline) as pool:
            list(pool.map(lane_loop, lanes))
        barrier()
        # wall time of e2e_steps steps of EVERY lane -> time per e2e_steps steps (the serial figure's convention)
        e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3) / e2e_pipeline
        for d2, s2, _ in lanes[1:]:
            s2.close()
            d2.close()
    e2e_value = (float(n * e2e_steps) if sharded else sum_over_ranks(float(n * e2e_steps))) / (e2e_ms * 1e-3)
    e2e_serial_value = (float(n * e2e_steps) if sharded else sum_over_ranks(float(n * e2e_steps))) / (e2e_serial_ms * 1e-3)
    comm_stats = dev.comm_stats() if sharded else None

    # ---- roofline of the dominant kernel (k_sweep, 8 launches per step) ----
    k_cand, k_full = dev.neighbour_stats()
    w_move = k_cand * F_PRE + k_full * F_FULL
    peak = dev.measure_fp64_peak()
    launches_sweep = 8 * args.steps
    avg_launch_ms = sweep_ms / launches_sweep
    units_per_launch = attempted / launches_sweep / (world if sharded else 1)    # per GPU
    achieved_tflops = units_per_launch * w_move / (avg_launch_ms * 1e-3) * 1e-12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("k_sweep_bytes_per_launch")
    except (OSError, ValueError):
        pass
    # the physical ceiling of the kernel as built: the FP32 capsule cull visits every (trial, staged candidate) pair of
    # a cell; its inner loop is CULL_INSTR_PER_PAIR issue slots per pair and lane
    cells = sweep_cells(np.array(b_now), n, 6.0)
    per_cell = n / float(cells[0] * cells[1] * cells[2])
    cull_pairs_per_move = 28.0 * per_cell
    sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
    cull_peak = 148 * 4 * 32 * sm_mhz * 1e6 / CULL_INSTR_PER_PAIR
    cull_rate = units_per_launch * cull_pairs_per_move / (avg_launch_ms * 1e-3)
    roofline = {"bound": "fp64", "kernel": "k_sweep<spherocylinder>", "achieved": achieved_tflops,
                "peak": peak["dmul_dadd_gflops"] * 1e-3, "unit": "TFLOP/s",
                "frac": achieved_tflops / (peak["dmul_dadd_gflops"] * 1e-3),
                "peak_source": "measured live: non-fused DMUL+DADD microbenchmark (rp_measure_fp64_peak); "
                               f"fused DFMA peak {peak['dfma_gflops'] * 1e-3:.2f} TFLOP/s",
                "flop_per_move": w_move, "k_cand": k_cand, "k_full": k_full, "avg_launch_ms": avg_launch_ms,
                "kernel_share_of_step": sweep_ms / ms if world == 1 else None,
                "traffic": traffic,
                "physical": {"what": "FP32 capsule-cull pairs (every trial of an active cell against every particle "
                                     "staged for it) against the issue-slot ceiling of that loop",
                             "cull_pairs_per_move": cull_pairs_per_move, "achieved": cull_rate,
                             "peak": cull_peak, "unit": "pairs/s", "frac": cull_rate / cull_peak,
                             "peak_source": f"148 SMs x 4 schedulers x 32 lanes x {sm_mhz:.0f} MHz / "
                                            f"{CULL_INSTR_PER_PAIR:g} issue slots per pair (SASS of the inner loop)"},
                "hbm": {"achieved": units_per_launch * B_MOVE / (avg_launch_ms * 1e-3) * 1e-9, "peak": hbm_peak,
                        "unit": "GB/s", "frac": units_per_launch * B_MOVE / (avg_launch_ms * 1e-3) * 1e-9 / hbm_peak,
                        "bytes_per_move": B_MOVE,
                        "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback"}}

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cm, cs, csec, info, _ = run_reference_cycles(therm_box, therm_shapes, args.cpu_cycles, SEED, steps=steps3)
            cpu_baseline = {"value": cm / csec, "unit": UNIT, "cores": info["cores"], "kind": info["kind"],
                            "sample": f"{args.cpu_cycles} NpT cycle(s) of the same {args.workload} state the GPU "
                                      f"timing starts from (N={n}, all particles), domains {info['domains']}",
                            "flags": info["flags"], "move_seconds": info["move_seconds"],
                            "scaling_seconds": info["scaling_seconds"]}
        except Exception as ex:      # the baseline is a reported figure; its absence must not hide the GPU number
            cpu_baseline = {"value": None, "unit": UNIT, "cores": 0, "kind": "unavailable", "sample": str(ex)}
    sim.close()
    dev.close()

    # ---- N > 1, sharded: the other way north_star uses several GPUs — independent pressure points of a scan, one
    # packing per GPU, no data-path collective — measured in the same run and reported as an extra key ----
    replicas = None
    if sharded:
        rdev, rsim, rsteps = thermalised_state(args.workload, args.therm, local, SEED + rank, PRESSURE + 0.1 * rank,
                                               stream.cuda_stream, None)
        rsim.run(args.warmup)
        barrier()
        with torch.cuda.stream(stream):
            e0.record(stream)
            rsim.run(args.steps)
            e1.record(stream)
        barrier()
        rms = max_over_ranks(e0.elapsed_time(e1))
        replicas = {"value": sum_over_ranks(float(n * args.steps)) / (rms * 1e-3), "unit": UNIT, "scaling": "weak",
                    "ms_per_step": rms / args.steps,
                    "what": f"{world} independent {args.workload} packings (p = 1.2 + 0.1 rank), one per GPU"}
        rsim.close()
        rdev.close()

    # ---- the other BASELINE configurations (C1 - C4), N = 1 only: extra keys of the same line ----
    configs = None
    if rank == 0 and world == 1 and args.configs != "none":
        names = list(W.EXTRA_CONFIGS) if args.configs == "all" else [x for x in args.configs.split(",") if x]
        configs = {}
        for name in names:
            try:
                configs[name] = measure_extra(name, W.EXTRA_CONFIGS[name], local, stream, peak["dmul_dadd_gflops"],
                                              args.steps, cpu=not args.no_cpu_baseline)
            except Exception as ex:
                configs[name] = {"value": None, "error": str(ex)}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "strong" if sharded else "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload_string(args.workload, n, args.therm, steps3, sharded, world > 1),
                           "parallelism": ("1 GPU" if world == 1 else
                                           f"spatially sharded sweep: {world} z-slabs of cell layers; accepted moves "
                                           "travel as accept bytes + boundary-layer records written by k_sweep into "
                                           "the peers' memory over NVLink, flag kernels order the colours; box-move "
                                           "recount split + NCCL all-reduce; state replicated"
                                           if sharded else "replicas (pressure-scan points), one per GPU"),
                           "l2_policy": f"state {n * 96 / 1e6:.0f} MB + sorted records {n * 80 / 1e6:.0f} MB per "
                                        "step; every step rebuilds the sweep grid from all particles "
                                        "(inputs not flushed: working set is re-read each step by design)"},
                "e2e": {"value": e2e_value, "unit": UNIT,
                        "h2d_bytes_per_step": int(n * 96 + 72 * world) if sharded else int((n * 96 + 72) * world),
                        "d2h_bytes_per_step": int(n * 96 + 72 * world) if sharded else int((n * 96 + 72) * world),
                        "steps": e2e_steps * e2e_pipeline, "ms_per_step": e2e_ms / e2e_steps,
                        "packings_in_flight_per_gpu": e2e_pipeline,
                        "serial_value": e2e_serial_value, "serial_ms_per_step": e2e_serial_ms / e2e_steps,
                        "note": ("every rank moves its slice of the one packing (upload + all-gather over NVLink, "
                                 "cycle, download of the slice), one packing, nothing overlapped" if sharded
                                 else f"every step = upload of the whole packing from pinned host memory + one NpT "
                                      f"cycle + download of the whole packing; {e2e_pipeline} independent packings "
                                      "per GPU in flight on separate handles / streams so that copies overlap "
                                      "compute; serial_value = one packing, nothing overlapped")},
                "gpu_launches": int(launches),
                "clocks": clocks,
                "roofline": roofline,
                "cpu_baseline": cpu_baseline,
                "acceptance": accepted / max(1, attempted),
                "step_sizes": {"translation": steps3[0], "rotation": steps3[1], "box": steps3[2]},
                "volume_moves": {"attempted": args.steps, "evaluated": int(evaluated),
                                 "accepted": int(scalings_accepted),
                                 "evaluated_per_s": (evaluated / (scaling_ms * 1e-3)) if scaling_ms > 0 else None,
                                 "ms_per_evaluated": (scaling_ms / evaluated) if evaluated else None},
                "comm_bytes_last_sweep": comm_stats,
                "breakdown_ms_per_step": {"sweep_kernels": sweep_ms / args.steps, "sweep_grid": grid_ms / args.steps,
                                          "box_move": scaling_ms / args.steps},
                "replicas": replicas,
                "configs": configs}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
