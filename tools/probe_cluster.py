"""GPU probe: cluster stage of single C2 frames and of the batch, shared-memory path vs L2 path."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200"))
sys.path.insert(0, ROOT)
import numpy as np
import pcseg, bench
frames = bench.gen_frames(list(range(256)))
ctx = pcseg.Context(0)
params = pcseg.pipeline_params(seed_mode=pcseg.SEEDS_ALL_DESCENDING, run_cluster=True)
b = pcseg.Batch(ctx, 256, 480, 640)
b.run(frames, params)
st = b.frame_stats()
big = np.argsort(-st[:, 3])[:3]
print("largest cluster inliers:", [(int(i), int(st[i, 3]), int(st[i, 1])) for i in big], "mean cluster inliers", float(st[:, 3].mean()))
b.close()
for mode in ("wave", "global"):
    if mode == "global":
        os.environ["PCSEG_CLUSTER_MODE"] = "global"
    out = {}
    for name, ids in (("big0", [int(big[0])]), ("frame0", [0]), ("all", list(range(256)))):
        bb = pcseg.Batch(ctx, len(ids), 480, 640)
        for _ in range(3):
            bb.run(frames[ids], params)
        s = bb.frame_stats()
        out[name] = {"cluster_ms": round(bb.stage_times()["cluster"], 3), "steps_incl_planar": int(s[:, 6].max())}
        bb.close()
    print(mode, json.dumps(out))
