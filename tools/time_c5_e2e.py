"""GPU probe: SlidingMeanShift on the full-size C5 cloud through the public call with HOST arrays
(PCSEG_MS_DEBUG=1 prints the phases)."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200"))
import numpy as np
import pcseg, synth
rows, cols = 2500, 4000
xyz = synth.c5_full(rows, cols)
labels = np.full(rows * cols, -1, np.int32)
ctx = pcseg.Context(0)
ts = []
for rep in range(3):
    t0 = time.perf_counter()
    lab, r = ctx.mean_shift(xyz, rows, cols, labels, iterations=5)
    ts.append((time.perf_counter() - t0) * 1e3)
    sys.stderr.write("---- rep %d: %.1f ms, %d regions\n" % (rep, ts[-1], r.n))
print(json.dumps({"e2e_ms": [round(t, 1) for t in ts], "regions": int(r.n)}))
