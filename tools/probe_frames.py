"""GPU probe: per-frame work and how the growth stages scale with the number of
resident frames (what bounds the batch: the slowest frame or SM sharing)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200"))
sys.path.insert(0, ROOT)
import numpy as np
import pcseg, bench

frames = bench.gen_frames(list(range(256)))
ctx = pcseg.Context(0)
params = pcseg.pipeline_params(seed_mode=pcseg.SEEDS_ALL_DESCENDING, run_cluster=True)
res = {}
b = pcseg.Batch(ctx, 256, 480, 640)
nrm = np.empty((256, 480 * 640, 3), np.float32); lab = np.empty((256, 480 * 640), np.int32)
for _ in range(3):
    b.run(frames, params, normals_out=nrm, labels_out=lab)
res["t256"] = b.stage_times()
st = b.frame_stats()
res["stats_cols"] = "n_planar n_cluster planar_inl cluster_inl boundary err steps refits"
res["stats_sample"] = st[::16].tolist()
res["steps_minmax"] = [int(st[:, 6].min()), int(st[:, 6].max())]
res["planar_inl_minmax"] = [int(st[:, 2].min()), int(st[:, 2].max())]
res["nan_frac"] = float(np.isnan(frames[:, :, 0]).mean())
b.close()
for F in (1, 16, 74, 148, 222):
    bb = pcseg.Batch(ctx, F, 480, 640)
    for _ in range(3):
        bb.run(frames[:F], params)
    res["t%d" % F] = bb.stage_times()
    bb.close()
single = {}
b1 = pcseg.Batch(ctx, 1, 480, 640)
for i in range(0, 256, 32):
    for _ in range(2):
        b1.run(frames[i:i + 1], params)
    single[i] = b1.stage_times()
res["single"] = single
print(json.dumps(res))
