"""Scratch probe (not a test): stage timings of the batch pipeline on the GPU."""
import sys, time, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200"), os.path.join(ROOT, "oracle")]
import numpy as np
import pcseg, synth, oracle
F = int(sys.argv[1]) if len(sys.argv) > 1 else 16
rows, cols = 480, 640
t = time.time()
frames = synth.c2_batch(F, rows, cols, distinct=min(F, 8))
print("gen %.1fs" % (time.time() - t), flush=True)
ctx = pcseg.Context(0)
b = pcseg.Batch(ctx, F, rows, cols)
for mode, name in ((pcseg.SEEDS_ALL_DESCENDING, "alldesc"), (pcseg.SEEDS_PLANE_SUPPORT, "planesupport")):
    p = pcseg.pipeline_params(seed_mode=mode)
    labels = np.zeros((F, rows * cols), np.int32)
    for it in range(3):
        t = time.time(); b.run(frames, p, labels_out=labels); dt = time.time() - t
        print(name, "run %d: %.1f ms  (%.1f Mpts/s)" % (it, dt * 1e3, F * rows * cols / dt / 1e6), flush=True)
    print(name, "stage ms", b.stage_times(), flush=True)
    try:
        pr = b.profile()
        tot = pr.sum(axis=0).astype(float)
        print(name, "phase cycles/step: eval %.0f cut+accum %.0f nbr+claim %.0f enqueue %.0f refit %.0f | cut+stores %.0f masktest+rej %.0f" % tuple(tot[:7] / max(b.frame_stats()[:, 6].sum(), 1) / 3))
    except Exception as e:
        print("no profile", e)
    st = b.frame_stats()
    print("stats frame0", st[0], "max steps", st[:, 6].max(), "err", st[:, 5].max())
    # parity of frame 0 and last frame vs oracle
    for f in (0, F - 1):
        xyz = frames[f]
        o_n = oracle.normals_organized(xyz, rows, cols)
        if mode == pcseg.SEEDS_ALL_DESCENDING:
            seeds = np.arange(rows * cols - 1, -1, -1, dtype=np.int32)
        else:
            seeds = oracle.seeds_plane_support(xyz, o_n, rows, cols)
        lab0 = np.full(rows * cols, -1, np.int32)
        o_lab, o_r = oracle.segment_planar(xyz, o_n, rows, cols, seeds, lab0)
        allseeds = np.arange(rows * cols - 1, -1, -1, dtype=np.int32)
        o_lab2, o_c = oracle.segment_cluster(xyz, rows, cols, allseeds, o_lab, id_offset=o_r.n)
        print(name, "frame", f, "labels equal:", np.array_equal(labels[f], o_lab2), "planes", o_r.n, "clusters", o_c.n, flush=True)
