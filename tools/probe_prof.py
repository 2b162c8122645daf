"""GPU probe (PCSEG_PROFILE build, `make prof`): per-role cycle breakdown of planar growth."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ["PCSEG_LIB"] = os.path.join(ROOT, "build", "libpcseg_prof.so")
sys.path.insert(0, os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200"))
sys.path.insert(0, ROOT)
import numpy as np
import pcseg, bench
frames = bench.gen_frames(list(range(256)))
ctx = pcseg.Context(0)
params = pcseg.pipeline_params(seed_mode=pcseg.SEEDS_ALL_DESCENDING, run_cluster=True)
names = ["A.refit_apply", "A.verify", "A.spec", "A.idle", "A.qlen_sum", "A.rollbacks", "A.specdepth_sum", "cluster_cycles",
         "B.setup", "B.claims", "B.enqueue", "B.wait", "B.passes", "seed_scan", "finalize", "planar_cycles"]
for F in (16, 128, 256):
    b = pcseg.Batch(ctx, F, 480, 640)
    for _ in range(2):
        b.run(frames[:F], params)
    b.profile()
    b.run(frames[:F], params)
    pr = b.profile().astype(np.float64)
    st = b.frame_stats()
    t = b.stage_times()
    print("F=%d planar %.2f ms cluster %.2f ms; steps/frame %.0f refits %.0f" % (F, t["planar"], t["cluster"], st[:, 6].mean(), st[:, 7].mean()))
    m = pr.mean(axis=0)
    for n, v in zip(names, m):
        if True:
            print("   %-18s %12.0f  (per step %.1f)" % (n, v, v / st[:, 6].mean()))
    tot = (pr[:, 0] + pr[:, 1] + pr[:, 2] + pr[:, 3] + pr[:, 13] + pr[:, 14]) / 1.965e6
    print("   per-frame total ms (A+seed+fin): mean %.2f min %.2f max %.2f p90 %.2f" % (tot.mean(), tot.min(), tot.max(), np.percentile(tot, 90)))
    pl, cl = pr[:, 15] / 1.965e6, pr[:, 7] / 1.965e6
    print("   planar ms/frame: mean %.2f min %.2f max %.2f | cluster: mean %.2f max %.2f | sum: max %.2f" % (pl.mean(), pl.min(), pl.max(), cl.mean(), cl.max(), (pl + cl).max()))
    o = np.argsort(-(pl + cl))[:8]
    print("   slowest frames:", [(int(i), round(float(pl[i]), 1), round(float(cl[i]), 1), int(st[i, 2]), int(st[i, 3])) for i in o])
    b.close()
