"""Runs the C2 pipeline on F frames three times (for ncu captures of single launches)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200"))
sys.path.insert(0, ROOT)
import torch
import pcseg, bench
F = int(sys.argv[1]) if len(sys.argv) > 1 else 16
frames = bench.gen_frames(list(range(F)))
ctx = pcseg.Context(0)
params = pcseg.pipeline_params(seed_mode=pcseg.SEEDS_ALL_DESCENDING, run_cluster=True)
b = pcseg.Batch(ctx, F, 480, 640)
dev = torch.from_numpy(frames).cuda()   # device-resident input: one launch per stage for the whole batch
for _ in range(3):
    b.run(dev.data_ptr(), params)
print(b.stage_times())
