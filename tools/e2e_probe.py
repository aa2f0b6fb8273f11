"""Development probe: where does the time go when several packings are in flight on one GPU (upload / cycle / download
per handle, one host thread each)?"""
import os, sys, time, threading
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import rampack_b200 as rp
from rampack_b200 import workloads as W, NativeSimulation

lanes_n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
mode = sys.argv[2] if len(sys.argv) > 2 else "ucd"       # which phases run: u(pload) c(ycle) d(ownload)
steps = 8
box, shapes = W.sc_config("sc_1m")
n = shapes.shape[0]
lanes = []
for k in range(lanes_n):
    d = rp.Device(0, seed=10 + k)
    d.set_interaction(rp.make_desc(hard=rp.RP_HARD_SPHEROCYLINDER, radius=0.5, length=5.0))
    d.upload(box, shapes)
    s = NativeSimulation(d, [(rp.RP_MOVE_ROTOTRANSLATION, 0.05, 0.05)], 1.0, 1.2, ("delta_triclinic", 1e-5, False), seed=k)
    s.run(3)
    h = torch.empty((n, 12), dtype=torch.float64).pin_memory().numpy()
    h[:] = shapes
    lanes.append((d, s, h))
times = [[] for _ in lanes]
outs = [[torch.empty((n, 12), dtype=torch.float64).pin_memory().numpy() for _ in range(2)] for _ in lanes]
def loop(i):
    d, s, h = lanes[i]
    bx = box
    for _ in range(steps):
        t0 = time.perf_counter()
        if "u" in mode: d.upload(bx, h)
        t1 = time.perf_counter()
        if "c" in mode: s.run(1)
        t2 = time.perf_counter()
        if "d" in mode: bx, _ = d.download(out=h)
        if "s" in mode:            # asynchronous snapshot into a second pinned buffer (D2H on the snapshot stream)
            d.snapshot_wait()      # the previous snapshot must have left the device-side copy / its host buffer
            bx = d.snapshot_async(outs[i][len(times[i]) % 2])
        t3 = time.perf_counter()
        times[i].append((t1 - t0, t2 - t1, t3 - t2))
    if "s" in mode: d.snapshot_wait()
torch.cuda.synchronize()
t0 = time.perf_counter()
th = [threading.Thread(target=loop, args=(i,)) for i in range(lanes_n)]
[t.start() for t in th]; [t.join() for t in th]
wall = time.perf_counter() - t0
print(f"mode {mode} lanes {lanes_n}: wall {wall*1e3:.1f} ms for {steps*lanes_n} steps = {wall*1e3/(steps*lanes_n):.2f} ms/step")
for i, t in enumerate(times):
    a = np.array(t[2:]) * 1e3
    print(f"  lane {i}: upload {a[:,0].mean():.2f}  cycle {a[:,1].mean():.2f}  download {a[:,2].mean():.2f} ms")
