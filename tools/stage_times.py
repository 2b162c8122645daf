"""GPU probe: stage times of one 256-frame C2 batch alone (and of 1 / 16 frames) for the library
named by PCSEG_LIB (default: the in-tree build); checks labels of 4 frames against the oracle."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, ROOT)
import numpy as np
import pcseg, bench
frames = bench.gen_frames(list(range(256)))
ctx = pcseg.Context(0)
params = pcseg.pipeline_params(seed_mode=pcseg.SEEDS_ALL_DESCENDING, run_cluster=True)
out = {"lib": os.environ.get("PCSEG_LIB", "default")}
for F in (1, 16, 256):
    b = pcseg.Batch(ctx, F, 480, 640)
    lab = np.empty((F, 480 * 640), np.int32)
    for _ in range(3):
        b.run(frames[:F], params, labels_out=lab)
    out["F%d" % F] = {k: round(v, 2) for k, v in b.stage_times().items()}
    if F == 16 and "--check" in sys.argv:
        bad = 0
        for i in (0, 5, 10, 15):
            l2, _, _ = bench._cpu_pipeline(frames[i], want_regions=True)
            bad += int(not np.array_equal(l2, lab[i]))
        out["label_mismatch_frames_of_4"] = bad
    b.close()
print(json.dumps(out))
