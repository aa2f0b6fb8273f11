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
TRANS_STEP, ROT_STEP = 0.05, 0.05      # rototranslation(trans_step=0.05, rot_step=0.05)
TEMPERATURE, PRESSURE = 1.0, 1.2
SEED = 1234
# flops of one pair test in the reference arithmetic: centre-distance pre-filter, full segment test
F_PRE, F_FULL = 14, 105
B_MOVE = 200                           # algorithmic bytes per attempted move: state read + write-back + list entry


def box_step_for(n):
    """delta_triclinic step at which N ln(V'/V) is O(1) — where the reference's step adaptation settles
    (Simulation.cpp:533-607); 0.01 at N = 1e5 (Appendix C) scaled with the particle number."""
    return 0.01 * 1.0e5 / max(n, 1.0e5) if n > 1.0e5 else 0.01


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


def run_reference_cycles(box, shapes, cycles, seed, threads=None):
    """`cycles` NpT cycles of the reference's own Simulation (oracle/_ref when built, else the single-thread port).
    Returns (moves attempted, scalings attempted, seconds for moves + scaling + domain decomposition, info, state)."""
    from oracle import refwrap as R
    from rampack_b200 import workloads as W
    import rampack_b200 as rp
    desc = rp.make_desc(hard=rp.RP_HARD_SPHEROCYLINDER, radius=W.SC_RADIUS, length=W.SC_LENGTH)
    n = shapes.shape[0]
    if R.have_ref(native=True) or R.have_ref():
        native = R.have_ref(native=True)
        traits = R.RefTraits(desc, native=native)
        cores = threads or host_cores()
        heights = [box[0, 0], box[1, 1], box[2, 2]]
        dom = choose_domains(cores, heights, 6.0, 6.2)
        nthreads = dom[0] * dom[1] * dom[2]
        res, box_out, shapes_out = R.ref_run_simulation(traits, box, shapes, seed, TEMPERATURE, PRESSURE, TRANS_STEP,
                                                        ROT_STEP, 2, box_step_for(n), dom, nthreads, cycles, 0)
        seconds = res.move_seconds + res.scaling_seconds + res.domain_decomposition_seconds
        info = {"kind": "reference", "cores": nthreads, "domains": list(dom),
                "flags": "-O3 -march=x86-64-v3 -fopenmp" if native else "-O3 -ffp-contract=off -fopenmp",
                "move_seconds": res.move_seconds, "scaling_seconds": res.scaling_seconds,
                "scalings_attempted": int(res.scalings_attempted)}
        return int(res.moves_attempted), int(res.scalings_attempted), seconds, info, (box_out, shapes_out)
    pk = R.PortPacking(R.PortTraits(desc), box, shapes)
    res = pk.run_cycles(seed, cycles, TRANS_STEP, ROT_STEP, TEMPERATURE, PRESSURE, True, box_step_for(n))
    info = {"kind": "port", "cores": 1, "domains": [1, 1, 1], "flags": "-O2 -ffp-contract=off",
            "move_seconds": res.move_seconds, "scaling_seconds": res.scaling_seconds,
            "scalings_attempted": int(res.scalings_attempted)}
    b, s = pk.get()
    return int(res.moves_attempted), int(res.scalings_attempted), res.move_seconds + res.scaling_seconds, info, (b, s)


def reference_workload(args):
    """The reference arm runs the metric's configuration when K + W cycles of it fit a few minutes on this host,
    otherwise the same lattice at N = 100 352 (same density, same per-move work)."""
    from rampack_b200 import workloads as W
    cores = host_cores()
    name = args.workload
    box, shapes = W.sc_config(name)
    est = shapes.shape[0] / (3.0e4 * max(1, min(cores, 512))) * (args.steps + args.warmup)
    if est > 150 and name != "sc_100k":
        name = "sc_100k"
        box, shapes = W.sc_config(name)
    return name, box, shapes


def main_reference(args):
    rank = env_int("RANK", 0)
    if rank != 0:
        return 0
    name, box, shapes = reference_workload(args)
    n = shapes.shape[0]
    state = (box, shapes)
    if args.warmup > 0:
        _, _, _, _, state = run_reference_cycles(box, shapes, args.warmup, SEED)
    moves, scalings, seconds, info, _ = run_reference_cycles(state[0], state[1], args.steps, SEED + 1)
    value = moves / seconds
    sample = f"{args.steps} NpT cycles of {name} (N={n}) after {args.warmup} warm-up cycles, all particles"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": seconds / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"{name}: hard spherocylinders l=5 r=0.5, N={n}, sc lattice eta=0.475, NpT "
                                   f"T=1 p=1.2, rototranslation(0.05, 0.05), delta_triclinic({box_step_for(n):g})",
                       "parallelism": f"OpenMP domains {info['domains']}"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": info["cores"], "kind": info["kind"],
                             "sample": sample, "flags": info["flags"]},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "volume_moves_per_s": (scalings / info["scaling_seconds"]) if info["scaling_seconds"] > 0 else None,
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="sc_1m", choices=["sc_12k", "sc_100k", "sc_1m", "sc_4m"])
    ap.add_argument("--therm", type=int, default=20, help="untimed equilibration cycles before the warm-up")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-cycles", type=int, default=1)
    ap.add_argument("--parallelism", default="sharded", choices=["sharded", "replicas"],
                    help="what N > 1 ranks do (ignored at N = 1)")
    args = ap.parse_args()
    if args.impl == "reference":
        return main_reference(args)

    import torch
    import rampack_b200 as rp
    from rampack_b200 import workloads as W
    from rampack_b200 import NativeSimulation

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
    dev = rp.Device(device=local, seed=rseed, cuda_stream=stream.cuda_stream)
    if sharded:
        ident = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            ident.copy_(torch.frombuffer(bytearray(rp.Device.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(ident, 0)
        dev.comm_init(bytes(ident.cpu().numpy().tobytes()), rank, world)
    desc = rp.make_desc(hard=rp.RP_HARD_SPHEROCYLINDER, radius=W.SC_RADIUS, length=W.SC_LENGTH)
    box, shapes = W.sc_config(args.workload)
    n = shapes.shape[0]
    dev.set_interaction(desc)
    dev.upload(box, shapes)
    # each replica is one pressure point of a scan (north_star): rank r runs p = 1.2 + 0.1 r
    # the reference's Simulation cycle loop, in C++ inside the library (rp_sim, rampack_b200/csrc/rp_sim.cu); step
    # sizes stay frozen (adjust_step_sizes = 0) so that both arms do the same work
    sim = NativeSimulation(dev, [(rp.RP_MOVE_ROTOTRANSLATION, TRANS_STEP, ROT_STEP)], TEMPERATURE,
                           PRESSURE if sharded else PRESSURE + 0.1 * rank,
                           ("delta_triclinic", box_step_for(n), False), seed=rseed)
    sim.run(args.therm)
    therm_box, therm_shapes = dev.download()

    # ---- device-resident timing: W warm-up + exactly K timed steps ----
    sim.run(args.warmup)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
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
    clocks = sampler.stop() if rank == 0 else None
    launches = dev.launch_count() - launches0
    s1 = sim.stats()
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
    b_now, s_now = dev.download()
    host = torch.empty((n, 12), dtype=torch.float64).pin_memory().numpy()
    host[:] = s_now
    e2e_box = b_now
    e2e_steps = max(3, min(args.steps, 10))
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        dev.upload(e2e_box, host)
        sim.run(1)
        e2e_box, _ = dev.download(out=host)
    barrier()
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
    e2e_value = (float(n * e2e_steps) if sharded else sum_over_ranks(float(n * e2e_steps))) / (e2e_ms * 1e-3)
    comm = dev.comm_stats() if sharded else None

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
    roofline = {"bound": "fp64", "kernel": "k_sweep<spherocylinder>", "achieved": achieved_tflops,
                "peak": peak["dmul_dadd_gflops"] * 1e-3, "unit": "TFLOP/s",
                "frac": achieved_tflops / (peak["dmul_dadd_gflops"] * 1e-3),
                "peak_source": "measured live: non-fused DMUL+DADD microbenchmark (rp_measure_fp64_peak); "
                               f"fused DFMA peak {peak['dfma_gflops'] * 1e-3:.2f} TFLOP/s",
                "flop_per_move": w_move, "k_cand": k_cand, "k_full": k_full, "avg_launch_ms": avg_launch_ms,
                "kernel_share_of_step": sweep_ms / ms if world == 1 else None,
                "traffic": traffic,
                "hbm": {"achieved": units_per_launch * B_MOVE / (avg_launch_ms * 1e-3) * 1e-9, "peak": hbm_peak,
                        "unit": "GB/s", "frac": units_per_launch * B_MOVE / (avg_launch_ms * 1e-3) * 1e-9 / hbm_peak,
                        "bytes_per_move": B_MOVE,
                        "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback"}}

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cm, cs, csec, info, _ = run_reference_cycles(therm_box, therm_shapes, args.cpu_cycles, SEED)
            cpu_baseline = {"value": cm / csec, "unit": UNIT, "cores": info["cores"], "kind": info["kind"],
                            "sample": f"{args.cpu_cycles} NpT cycle(s) of the same {args.workload} state the GPU "
                                      f"timing starts from (N={n}, all particles), domains {info['domains']}",
                            "flags": info["flags"], "move_seconds": info["move_seconds"],
                            "scaling_seconds": info["scaling_seconds"]}
        except Exception as ex:      # the baseline is a reported figure; its absence must not hide the GPU number
            cpu_baseline = {"value": None, "unit": UNIT, "cores": 0, "kind": "unavailable", "sample": str(ex)}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "strong" if sharded else "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": f"{args.workload}: hard spherocylinders l=5 r=0.5, N={n} "
                                       f"{'in ONE packing sharded over all GPUs' if sharded else 'per GPU'}, sc lattice "
                                       f"eta=0.475 + {args.therm} equilibration cycles, NpT T=1 p=1.2, "
                                       f"rototranslation(0.05, 0.05), delta_triclinic({box_step_for(n):g})",
                           "parallelism": ("1 GPU" if world == 1 else
                                           f"spatially sharded sweep: {world} z-slabs of cell layers; accepted moves "
                                           "travel as accept bytes + boundary-layer records written by k_sweep into "
                                           "the peers' memory over NVLink, flag kernels order the colours; box-move "
                                           "recount split + NCCL all-reduce; state replicated"
                                           if sharded else "replicas (pressure-scan points), one per GPU"),
                           "l2_policy": f"state {n * 96 / 1e6:.0f} MB + sorted records {n * 48 / 1e6:.0f} MB per "
                                        "step; every step rebuilds the sweep grid from all particles "
                                        "(inputs not flushed: working set is re-read each step by design)"},
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(n * 96 + 72),
                        "d2h_bytes_per_step": int(n * 96 + 72), "steps": e2e_steps, "ms_per_step": e2e_ms / e2e_steps},
                "gpu_launches": int(launches),
                "clocks": clocks,
                "roofline": roofline,
                "cpu_baseline": cpu_baseline,
                "acceptance": accepted / max(1, attempted),
                "volume_moves": {"attempted": args.steps, "evaluated": int(evaluated),
                                 "accepted": int(scalings_accepted),
                                 "evaluated_per_s": (evaluated / (scaling_ms * 1e-3)) if scaling_ms > 0 else None,
                                 "ms_per_evaluated": (scaling_ms / evaluated) if evaluated else None},
                "comm_bytes_last_sweep": comm,
                "breakdown_ms_per_step": {"sweep_kernels": sweep_ms / args.steps, "sweep_grid": grid_ms / args.steps,
                                          "box_move": scaling_ms / args.steps}}
        print(json.dumps(line))
    dev.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
