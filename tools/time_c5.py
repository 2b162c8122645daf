"""GPU probe: phases of SlidingMeanShift on the full-size C5 cloud (device-resident input)."""
import ctypes as C, os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200"))
import numpy as np, torch
import pcseg, synth
rows, cols = 2500, 4000
n = rows * cols
xyz = synth.c5_full(rows, cols)
d_xyz = torch.from_numpy(xyz).cuda()
d_lab = torch.full((n,), -1, dtype=torch.int32, device="cuda")
ctx = pcseg.Context(0)
L = ctx.L
modes = torch.empty((n, 8), dtype=torch.float32, device="cuda")
def t(fn, reps=3):
    ts = []
    for _ in range(reps):
        torch.cuda.synchronize(); t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
    return round(min(ts), 2)
out = {}
out["modes_ms"] = t(lambda: ctx._check(L.pcseg_mean_shift_modes(ctx.h, C.c_void_p(d_xyz.data_ptr()), rows, cols, C.c_void_p(d_lab.data_ptr()), 5, C.c_void_p(modes.data_ptr()))))
inten = torch.where((modes[:, 7] != 0) & (modes[:, 6] != 0), modes[:, 5], torch.full_like(modes[:, 5], -1.0)).contiguous()
order = torch.empty((n,), dtype=torch.int32, device="cuda"); cnt = C.c_int(0)
out["sort_ms"] = t(lambda: ctx._check(L.pcseg_mean_shift_sort_order(ctx.h, C.c_void_p(inten.data_ptr()), n, 0, C.c_void_p(order.data_ptr()), C.byref(cnt))))
c = pcseg.cluster_config()
def from_modes():
    lab = d_lab.clone(); torch.cuda.synchronize(); h = C.c_void_p()
    l0 = ctx.launches
    ctx._check(L.pcseg_mean_shift_from_modes(ctx.h, C.c_void_p(d_xyz.data_ptr()), rows, cols, C.c_void_p(modes.data_ptr()), C.byref(c), 0, 0, C.c_void_p(lab.data_ptr()), C.byref(h)))
    out["regions"] = int(L.pcseg_regions_count(h)); out["launches_from_modes"] = ctx.launches - l0
    L.pcseg_regions_free(h)
out["sort_plus_growth_ms"] = t(from_modes)
out["growth_ms_by_difference"] = round(out["sort_plus_growth_ms"] - out["sort_ms"], 2)
print(json.dumps(out))
