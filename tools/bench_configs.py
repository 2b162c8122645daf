"""Secondary timings (not the bench line): configs C3, C4, C5 at full size on one GPU, device
time with CUDA events around the C-ABI calls.  Run through `python bench.py --secondary`, which
passes the CPU checker in for the cpu_baseline columns (this file itself never touches oracle/);
run directly it reports the GPU side only.  Prints one JSON object."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200"))
import numpy as np
import torch
import pcseg, synth

cpu_ref = globals().get("CPU_BASELINE_MODULE")  # injected by bench.py --secondary

ctx = pcseg.Context(0)
stream = torch.cuda.ExternalStream(ctx.stream)


def gpu_ms(fn, reps=5):
    fn(); fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(reps):
        fn()
    e1.record(stream)
    e1.synchronize()
    return e0.elapsed_time(e1) / reps


def cpu_ms(fn):
    t = time.perf_counter(); fn(); return (time.perf_counter() - t) * 1e3


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()

out = {}
# ---- C3 ----
rows, cols = 720, 1280
xyz = synth.c3(rows, cols); N = rows * cols
d_xyz = dev(xyz); d_n = torch.empty((N, 3), dtype=torch.float32, device="cuda")
seeds = np.arange(N - 1, -1, -1, dtype=np.int32)
import ctypes as C
L = ctx.L
VP = C.c_void_p
np_ = pcseg.normals_params()
st3 = (C.c_double * 3)(0.0, 0.0, 0.0)
t_n = gpu_ms(lambda: ctx._check(L.pcseg_normals_organized(ctx.h, VP(d_xyz.data_ptr()), rows, cols, st3, C.byref(np_),
                                                          0, rows, 0, cols, VP(d_n.data_ptr()), None)))
nrm = d_n.cpu().numpy()
labels = np.full(N, -1, np.int32)
# growth stages through the numpy API (host buffers: the copies are < 1 ms of these times)
t_p = gpu_ms(lambda: ctx.segment_planar(xyz, nrm, rows, cols, seeds, labels), reps=3)
g_lab, g_r = ctx.segment_planar(xyz, nrm, rows, cols, seeds, labels)
t_c = gpu_ms(lambda: ctx.segment_cluster(xyz, rows, cols, seeds, g_lab, id_offset=g_r.n), reps=3)
NA = float("nan")
c_n = cpu_ms(lambda: cpu_ref.normals_organized(xyz, rows, cols)) if cpu_ref else NA
c_p = cpu_ms(lambda: cpu_ref.segment_planar(xyz, nrm, rows, cols, seeds, labels)) if cpu_ref else NA
c_c = cpu_ms(lambda: cpu_ref.segment_cluster(xyz, rows, cols, seeds, g_lab, id_offset=g_r.n)) if cpu_ref else NA
out["C3_1280x720"] = {"points": N, "planes": int(g_r.n), "gpu_ms": {"normals": t_n, "planar": t_p, "cluster": t_c},
                      "cpu_port_ms": {"normals": c_n, "planar": c_p, "cluster": c_c},
                      "gpu_Mpts_s": N / (t_n + t_p + t_c) / 1e3, "cpu_Mpts_s": N / (c_n + c_p + c_c) / 1e3,
                      "normals_GBs": 24.0 * N / t_n / 1e6}
# ---- C4 ----
rows, cols = 128, 16384
xyz, labels = synth.c4(rows=rows, cols=cols, n_objects=200); N = rows * cols
d_xyz = dev(xyz); seeds = np.arange(N - 1, -1, -1, dtype=np.int32)
cfg = dict(squared_distance_threshold=0.25, half_search_window=2, min_region_inliers=10)
t_c = gpu_ms(lambda: ctx.segment_cluster(xyz, rows, cols, seeds, labels, cfg=cfg), reps=3)
c_c = cpu_ms(lambda: cpu_ref.segment_cluster(xyz, rows, cols, seeds, labels, cfg=cpu_ref.default_cluster_config(**cfg))) if cpu_ref else NA
out["C4_lidar_2M"] = {"points": N, "gpu_ms": {"cluster": t_c}, "cpu_port_ms": {"cluster": c_c},
                      "gpu_Mpts_s": N / t_c / 1e3, "cpu_Mpts_s": N / c_c / 1e3}
# C4 extension variant: the same sweep shuffled into an unorganized cloud, voxel-hash clustering (radius 0.5 m)
perm = np.random.default_rng(3235).permutation(N)
u_xyz, u_lab = np.ascontiguousarray(xyz[perm]), np.ascontiguousarray(labels[perm])
d_u = dev(u_xyz); d_ul = dev(u_lab)


class _VCfg(C.Structure):
    _fields_ = [("radius", C.c_float), ("min_region_inliers", C.c_int)]


def _unorg():
    lab = d_ul.clone(); torch.cuda.synchronize(); h = VP(); cfgv = _VCfg(0.5, 10)
    ctx._check(L.pcseg_cluster_unorganized(ctx.h, VP(d_u.data_ptr()), N, C.byref(cfgv), 0, VP(lab.data_ptr()), C.byref(h)))
    n = L.pcseg_regions_count(h); L.pcseg_regions_free(h)
    return n
t_u = gpu_ms(_unorg, reps=3)
out["C4_unorganized_voxel_hash"] = {"points": N, "participating": int(((u_lab == -1) & np.isfinite(u_xyz[:, 0])).sum()),
                                    "regions": int(_unorg()), "gpu_ms": {"cluster_unorganized": t_u},
                                    "gpu_Mpts_s": N / t_u / 1e3,
                                    "cpu_port_ms": "not timed here: the sequential restatement needs ~80 s at this density"}
# ---- C5 ----
rows, cols = 2500, 4000
xyz = synth.c5_full(rows, cols); N = rows * cols
d_xyz = dev(xyz); labels = np.full(N, -1, np.int32); d_lab = dev(labels)
d_modes = torch.empty((N, 8), dtype=torch.float32, device="cuda")
t_m = gpu_ms(lambda: ctx._check(L.pcseg_mean_shift_modes(ctx.h, VP(d_xyz.data_ptr()), rows, cols, VP(d_lab.data_ptr()), 5,
                                                         VP(d_modes.data_ptr()))))
t_full = gpu_ms(lambda: ctx.mean_shift(xyz, rows, cols, labels, iterations=5), reps=2)
c_full = cpu_ms(lambda: cpu_ref.mean_shift(xyz, rows, cols, labels, iterations=5)) if cpu_ref else NA
n_seeds = int((~np.isnan(xyz[:, 0])).sum())
out["C5_meanshift_10M"] = {"points": N, "seeds": n_seeds, "gpu_ms": {"modes_kernel": t_m, "sliding_mean_shift_total": t_full},
                           "cpu_port_ms": {"sliding_mean_shift_total": c_full},
                           "modes_Gseeds_s": n_seeds / t_m / 1e6, "modes_GBs_algorithmic": 44.0 * n_seeds / t_m / 1e6,
                           "speedup_total": c_full / t_full}
print(json.dumps(out))
