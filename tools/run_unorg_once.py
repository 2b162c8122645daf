"""GPU probe: pcseg_cluster_unorganized on the shuffled C4 sweep (device-resident input), N calls.
For ncu launch lists (`ncu --metrics gpu__time_duration.sum -k regex:vox_|msort_|scan`)."""
import ctypes as C, os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200"))
import numpy as np, torch
import pcseg, synth
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
rows, cols = 128, 16384
xyz, labels = synth.c4(rows=rows, cols=cols, n_objects=200)
N = rows * cols
perm = np.random.default_rng(3235).permutation(N)
d_u = torch.from_numpy(np.ascontiguousarray(xyz[perm])).cuda()
d_ul = torch.from_numpy(np.ascontiguousarray(labels[perm])).cuda()
ctx = pcseg.Context(0)
L = ctx.L


class VCfg(C.Structure):
    _fields_ = [("radius", C.c_float), ("min_region_inliers", C.c_int)]


ts = []
for _ in range(reps):
    lab = d_ul.clone(); torch.cuda.synchronize(); h = C.c_void_p(); cfg = VCfg(0.5, 10)
    l0 = ctx.launches
    t0 = time.perf_counter()
    ctx._check(L.pcseg_cluster_unorganized(ctx.h, C.c_void_p(d_u.data_ptr()), N, C.byref(cfg), 0, C.c_void_p(lab.data_ptr()), C.byref(h)))
    ts.append((time.perf_counter() - t0) * 1e3)
    n = L.pcseg_regions_count(h); L.pcseg_regions_free(h)
print(json.dumps({"ms": [round(t, 2) for t in ts], "regions": int(n), "launches": int(ctx.launches - l0)}))
