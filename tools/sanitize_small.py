"""Small end-to-end pass of every entry point (for compute-sanitizer memcheck)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200"))
import numpy as np
import pcseg, synth
rows, cols = 96, 128
ctx = pcseg.Context(0)
xyz = synth.depth_frame(rows, cols, seed=5)
n = ctx.normals_organized(xyz, rows, cols, params=dict(min_neighbor_distance=0.05))
seeds = ctx.seeds_plane_support(xyz, n, rows, cols)
lab = ctx.prepare_unlabeled_labels(rows * cols)
l1, r1 = ctx.segment_planar(xyz, n, rows, cols, np.arange(rows * cols - 1, -1, -1, dtype=np.int32), lab)
l2, r2 = ctx.segment_cluster(xyz, rows, cols, np.arange(rows * cols - 1, -1, -1, dtype=np.int32), l1, id_offset=r1.n)
c5 = synth.c5(rows=90, cols=120, n_blobs=8, pitch=0.12)
l3, r3 = ctx.mean_shift(c5, 90, 120, np.full(90 * 120, -1, np.int32), iterations=3)
frames = np.stack([synth.depth_frame(rows, cols, seed=s) for s in range(4)])
b = [pcseg.Batch(ctx, 4, rows, cols) for _ in range(2)]
p = pcseg.pipeline_params(normals=pcseg.normals_params(min_neighbor_distance=0.05), run_cluster=True)
on = np.empty((4, rows * cols, 3), np.float32); ol = np.empty((4, rows * cols), np.int32)
b[0].submit(frames, p, normals_out=on, labels_out=ol); b[1].submit(frames, p)
b[0].wait(); b[1].wait()
print("ok", r1.n, r2.n, r3.n, len(seeds), b[1].frame_stats()[:, :2].tolist())
for x in b: x.close()
ctx.close()
