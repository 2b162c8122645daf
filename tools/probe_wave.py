"""GPU probe (PCSEG_PROFILE build, `make prof`): per-phase cycle breakdown of the window-parallel
planar growth (grow_planar_wave.cuh)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ["PCSEG_LIB"] = os.path.join(ROOT, "build", "libpcseg_prof.so")
sys.path.insert(0, os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200"))
sys.path.insert(0, ROOT)
import numpy as np
import pcseg, bench
frames = bench.gen_frames(list(range(256)))
ctx = pcseg.Context(0)
params = pcseg.pipeline_params(seed_mode=pcseg.SEEDS_ALL_DESCENDING, run_cluster=True)
names = ["verify(top)", "eval(->B1)", "scan+expand+commit", "leader.poll", "iteration", "rollbacks", "windows", "scan+stage+rec",
         "nbr tests+loads", "claims loop", "commit", "entries", "refits", "seed_scan", "finalize", "planar_cycles"]
for F in [int(x) for x in (sys.argv[1:] or ["1", "16", "256"])]:
    b = pcseg.Batch(ctx, F, 480, 640)
    for _ in range(2):
        b.run(frames[:F], params)
    b.profile()
    b.run(frames[:F], params)
    pr = b.profile().astype(np.float64)
    st = b.frame_stats()
    t = b.stage_times()
    print("F=%d planar %.2f ms cluster %.2f ms; windows/frame %.0f refits %.0f" % (F, t["planar"], t["cluster"], st[:, 6].mean(), st[:, 7].mean()))
    m = pr.mean(axis=0)
    w = max(m[6], 1)
    for n, v in zip(names, m):
        print("   %-18s %12.0f  (per window %.1f)" % (n, v, v / w))
    pl = pr[:, 15] / 1.965e6
    print("   planar ms/frame: mean %.2f min %.2f max %.2f" % (pl.mean(), pl.min(), pl.max()))
    b.close()
