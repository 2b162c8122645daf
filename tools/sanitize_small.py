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
# round-2 paths: mean-shift waves on touching regions (also with tiny inlier slots), one large frame through
# the connected-component cluster path, unorganized voxel clustering, the no-alloc seed finder
ter = synth.terraces(96, 120, 24, 30, 0.05, 0.6, seed=3)
for cap in (None, "100"):
    if cap: os.environ["PCSEG_MS_SLOT_CAP"] = cap
    l4, r4 = ctx.mean_shift(ter, 96, 120, np.full(96 * 120, -1, np.int32), cfg=dict(half_search_window=2), iterations=3, sort_mode=1)
    os.environ.pop("PCSEG_MS_SLOT_CAP", None)
spk = synth.speckle_cloud(512, 520, seed=9)
l5, r5 = ctx.segment_cluster(spk, 512, 520, np.arange(512 * 520 - 1, -1, -1, dtype=np.int32), np.full(512 * 520, -1, np.int32),
                             cfg=dict(squared_distance_threshold=0.004, half_search_window=1, min_region_inliers=7))
extra = (r4.n, r5.n)
rng = np.random.default_rng(4)
pts = np.concatenate([rng.normal(c, 0.15, size=(400, 3)) for c in ((0, 0, 0), (2, 0, 0), (0, 3, 1))]).astype(np.float32)
l6, r6 = ctx.cluster_unorganized(pts, np.full(len(pts), -1, np.int32), radius=0.3, min_region_inliers=5)
import torch
d_n = torch.from_numpy(np.ascontiguousarray(n, np.float32).reshape(-1)).cuda()
d_s = torch.zeros(rows * cols, dtype=torch.int32, device="cuda"); d_c = torch.zeros(1, dtype=torch.int32, device="cuda")
ctx.seeds_average_normals_noalloc(d_n.data_ptr(), rows, cols, d_s.data_ptr(), rows * cols, d_c.data_ptr(), 5, 8, 0.999)
print("ok", r1.n, r2.n, r3.n, len(seeds), b[1].frame_stats()[:, :2].tolist(), extra, r6.n, int(d_c.cpu()[0]))
for x in b: x.close()
ctx.close()
