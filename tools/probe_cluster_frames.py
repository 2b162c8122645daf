"""GPU probe: cluster-stage time of single C2 frames (one frame per batch), with the frame's statistics."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200"))
sys.path.insert(0, ROOT)
import numpy as np
import pcseg, bench
ids = [int(x) for x in sys.argv[1:]] or [7, 110, 117, 133, 200]
frames = bench.gen_frames(ids)
ctx = pcseg.Context(0)
params = pcseg.pipeline_params(seed_mode=pcseg.SEEDS_ALL_DESCENDING, run_cluster=True)
b = pcseg.Batch(ctx, 1, 480, 640)
for k, fid in enumerate(ids):
    f = frames[k:k + 1]
    for _ in range(3):
        b.run(f, params)
    t = b.stage_times(); st = b.frame_stats()[0]
    nan = int(np.isnan(f[0, :, 0]).sum())
    print(json.dumps({"frame": fid, "nan_cells": nan, "stage_ms": {k2: round(v, 3) for k2, v in t.items()}, "stats": [int(x) for x in st]}))
