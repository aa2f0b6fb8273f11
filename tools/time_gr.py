"""Times the device observables on the bench workload (sc_1m): nematic order, smectic order, g(r)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rampack_b200 as rp
from rampack_b200 import workloads as W

dev = rp.Device(0, seed=1)
dev.set_interaction(rp.make_desc(hard=rp.RP_HARD_SPHEROCYLINDER, radius=W.SC_RADIUS, length=W.SC_LENGTH))
box, shapes = W.sc_config(sys.argv[1] if len(sys.argv) > 1 else "sc_1m")
dev.upload(box, shapes)
for c in range(5):
    dev.sweep(rp.RP_MOVE_ROTOTRANSLATION, 0.2, 0.15, c)
for name, fn in (("nematic_order", lambda: dev.nematic_order()), ("smectic_order(5,5,5)", lambda: dev.smectic_order((5, 5, 5))),
                 ("pair_density_correlation(6.0, 60)", lambda: dev.pair_density_correlation(6.0, 60))):
    fn()
    t = time.time(); out = fn(); dt = time.time() - t
    print(f"{name}: {dt * 1e3:.2f} ms (N = {shapes.shape[0]})", flush=True)
counts, g = dev.pair_density_correlation(6.0, 60)
print("pairs inside 6.0:", int(counts.sum()), "g(r) tail:", g[-3:])
dev.close()
