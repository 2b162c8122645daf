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
        nch = max(b.frame_stats()[:, 6].sum(), 1) * 3
        print(name, "A cycles/chunk: finalize %.0f verify %.0f speculate %.0f idle-iters %.0f | per chunk: queue ahead %.1f rollbacks %.4f spec depth %.2f starved iters %.2f" % tuple(tot[:8] / nch))
        print(name, "B cycles/chunk: read+masktest %.0f claim %.0f enqueue+publish %.0f waiting %.0f | chunks %.2f" % tuple(tot[8:13] / nch))
        print(name, "kernel level, frame 0 per run: seed search %.2f ms, finalize %.2f ms, regions tried %d, below min inliers %d" % (pr[0, 13] / 3 / 1.965e6, pr[0, 14] / 3 / 1.965e6, pr[0, 15] // 3 // 1000000, (pr[0, 15] // 3) % 1000000))
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
