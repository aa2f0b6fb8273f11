"""Worker of tests/test_gpu_shard.py (one process per GPU, launched by torch.distributed.run): runs the same NpT
cycles sharded over all ranks and, on rank 0, on a single unsharded handle; the results must be bit-identical."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch
import torch.distributed as dist

import rampack_b200 as rp
import helpers as H


def run_cycles(dev, box, shapes, cycles, kind, part=None):
    dev.set_interaction(H.desc_sc())
    if part is None:
        dev.upload(box, shapes)
    else:
        # sharded upload: this rank sends its slice only, the ranks all-gather over NCCL (rp_upload_slice)
        first, count = rp.Device.slice_of(shapes.shape[0], part[1], part[0])
        dev.upload_slice(box, shapes[first:first + count], first, shapes.shape[0])
    log = []
    b = box.copy()
    for cycle in range(cycles):
        c = dev.sweep(kind, 0.25, 0.15, cycle)
        log.append((c.attempted, c.accepted, c.rejected_domain, c.rejected_overlap))
        # a compressing box move (usually rejected by an overlap) and an expanding one (accepted)
        for factor in (0.995, 1.002):
            nb = b * factor
            dE = dev.try_scaling(nb)
            log.append(dE)
            if dE == float("inf"):
                dev.revert_scaling()
            else:
                b = nb
        log.append(dev.count_total_overlaps(False))
    return log, dev.download()


def main():
    out_dir = sys.argv[1]
    kind = int(sys.argv[2]) if len(sys.argv) > 2 else rp.RP_MOVE_ROTOTRANSLATION
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ident = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        ident.copy_(torch.frombuffer(bytearray(rp.Device.comm_unique_id()), dtype=torch.uint8))
    dist.broadcast(ident, 0)
    box, shapes = H.sc_lattice((20, 20, 16))
    dev = rp.Device(device=local, seed=777)
    dev.comm_init(bytes(ident.cpu().numpy().tobytes()), rank, world)
    log, (b, s) = run_cycles(dev, box, shapes, 4, kind, part=(rank, world))
    first, count = rp.Device.slice_of(shapes.shape[0], world, rank)
    assert np.array_equal(dev.download_range(first, count), s[first:first + count])
    stats = dev.comm_stats()
    np.save(os.path.join(out_dir, f"state_{rank}.npy"), s)
    np.save(os.path.join(out_dir, f"box_{rank}.npy"), b)
    ok = True
    if rank == 0:
        single = rp.Device(device=local, seed=777)
        log1, (b1, s1) = run_cycles(single, box, shapes, 4, kind)
        ok = (log == log1) and np.array_equal(b, b1) and np.array_equal(s, s1)
        moved = int(np.any(s1 != shapes, axis=1).sum())
        print(f"RESULT ok={ok} moved={moved} accepted={[l[1] for l in log if isinstance(l, tuple)]} "
              f"halo_bytes={stats['halo_bytes']} gather_bytes={stats['gather_bytes']} "
              f"p2p={stats['halo_bytes'] == 0 and stats['gather_bytes'] == 0}", flush=True)
        if not ok:
            print("sharded log", log, "\nsingle  log", log1, flush=True)
        single.close()
    dist.barrier()
    dev.close()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
