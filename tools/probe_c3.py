"""GPU probe (PCSEG_PROFILE build, `make prof`): cycle breakdown of planar growth on the single
1280x720 frame of C3 (33 planes) -- where a frame that is one CTA spends its time."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ["PCSEG_LIB"] = os.path.join(ROOT, "build", "libpcseg_prof.so")
sys.path.insert(0, os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200"))
import numpy as np
import pcseg, synth
rows, cols = 720, 1280
xyz = synth.c3(rows, cols)
ctx = pcseg.Context(0)
params = pcseg.pipeline_params(seed_mode=pcseg.SEEDS_ALL_DESCENDING, run_cluster=True)
b = pcseg.Batch(ctx, 1, rows, cols)
frames = xyz.reshape(1, rows * cols, 3)
for _ in range(2):
    b.run(frames, params)
b.profile()
b.run(frames, params)
pr = b.profile().astype(np.float64)[0]
st = b.frame_stats()[0]
t = b.stage_times()
names = ["0 verify->B1 wait", "1 phase1 eval", "2", "3 leader poll", "4 iteration total", "5 rollbacks", "6 windows", "7 B1->scan",
         "8 scan+stage", "9 late claims+resolve", "10 append", "11 entries", "12 refits(J)", "13 seed_scan", "14 finalize", "15 planar_cycles"]
print(json.dumps({"stage_ms": t, "stats": [int(x) for x in st]}))
for n, v in zip(names, pr):
    print("  %-24s %14.0f  %8.3f ms" % (n, v, v / 1.965e6))
if pr[6] > 0:
    print("  entries per window %.1f, cycles per window %.0f" % (pr[11] / pr[6], pr[4] / pr[6]))
