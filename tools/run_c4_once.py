"""GPU probe: config C4 once through both clustering paths (device-resident where the ABI allows):
pcseg_cluster_unorganized on the shuffled sweep and pcseg_segment_cluster on the range image
(one CTA per connected component).  For ncu launch lists."""
import ctypes as C, os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200"))
import numpy as np, torch
import pcseg, synth
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
rows, cols = 128, 16384
xyz, labels = synth.c4(rows=rows, cols=cols, n_objects=200)
N = rows * cols
ctx = pcseg.Context(0)
L = ctx.L
perm = np.random.default_rng(3235).permutation(N)
d_u = torch.from_numpy(np.ascontiguousarray(xyz[perm])).cuda()
d_ul = torch.from_numpy(np.ascontiguousarray(labels[perm])).cuda()


class VCfg(C.Structure):
    _fields_ = [("radius", C.c_float), ("min_region_inliers", C.c_int)]


out = {"unorganized_ms": [], "organized_ms": []}
for _ in range(reps):
    lab = d_ul.clone(); torch.cuda.synchronize(); h = C.c_void_p(); cfg = VCfg(0.5, 10)
    t0 = time.perf_counter()
    ctx._check(L.pcseg_cluster_unorganized(ctx.h, C.c_void_p(d_u.data_ptr()), N, C.byref(cfg), 0, C.c_void_p(lab.data_ptr()), C.byref(h)))
    out["unorganized_ms"].append(round((time.perf_counter() - t0) * 1e3, 2))
    out["unorganized_regions"] = int(L.pcseg_regions_count(h)); L.pcseg_regions_free(h)
seeds = np.arange(N - 1, -1, -1, dtype=np.int32)
ccfg = dict(squared_distance_threshold=0.25, half_search_window=2, min_region_inliers=10)
for _ in range(reps):
    t0 = time.perf_counter()
    g_lab, g_r = ctx.segment_cluster(xyz, rows, cols, seeds, labels, cfg=ccfg)
    out["organized_ms"].append(round((time.perf_counter() - t0) * 1e3, 2))
    out["organized_regions"] = int(g_r.n)
print(json.dumps(out))
