"""ctypes binding of libpcseg_b200.so (the C ABI in include/pcseg_b200.h).

Thin host-side mirror used by tests/, bench.py and __graft_entry__.py.  There
is NO CPU fallback: if the CUDA library is missing or no GPU is present,
loading / Context() raises.
"""
import ctypes as C
import os
from types import SimpleNamespace

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PCSEG_LIB", os.path.join(_HERE, "libpcseg_b200.so"))
_LIB = None

F32P = C.POINTER(C.c_float)
I32P = C.POINTER(C.c_int)


class NormalsParams(C.Structure):
    _fields_ = [("min_neighbor_distance", C.c_float),
                ("max_neighbor_distance", C.c_float),
                ("include_diagonal_neighbors", C.c_int),
                ("min_num_support_neighbors", C.c_int)]


class PlanarConfig(C.Structure):
    _fields_ = [("max_plane_distance", C.c_float),
                ("min_region_area", C.c_float),
                ("min_region_inliers", C.c_int),
                ("plane_model_reestimation_period", C.c_int),
                ("discontinuity_min_range", C.c_float),
                ("discontinuity_max_range", C.c_float),
                ("discontinuity_normal_angle_diff", C.c_float),
                ("discontinuity_z_diff", C.c_float),
                ("discontinuity_z_ratio", C.c_float)]


class ClusterConfig(C.Structure):
    _fields_ = [("min_region_inliers", C.c_int),
                ("squared_distance_threshold", C.c_float),
                ("half_search_window", C.c_int)]


class PipelineParams(C.Structure):
    _fields_ = [("normals", NormalsParams), ("planar", PlanarConfig),
                ("cluster", ClusterConfig),
                ("sensor_translation", C.c_double * 3),
                ("robot_q_xyzw", C.c_float * 4),
                ("planar_seed_mode", C.c_int),
                ("planar_seeds", I32P), ("n_planar_seeds", C.c_int),
                ("run_cluster_stage", C.c_int)]


SEEDS_ALL_DESCENDING, SEEDS_PLANE_SUPPORT, SEEDS_EXPLICIT = 0, 1, 2

EXPORTS = """pcseg_default_normals_params pcseg_default_planar_config
pcseg_default_cluster_config pcseg_create pcseg_destroy pcseg_status_string
pcseg_last_error pcseg_launch_count pcseg_stream pcseg_prepare_unlabeled_labels
pcseg_normals_organized pcseg_segment_planar pcseg_segment_cluster
pcseg_mean_shift_modes pcseg_mean_shift_modes_range pcseg_mean_shift_from_modes
pcseg_mean_shift pcseg_seeds_plane_support
pcseg_batch_create pcseg_batch_destroy pcseg_batch_run pcseg_batch_submit pcseg_batch_wait pcseg_batch_stream pcseg_batch_frame_stats
pcseg_batch_regions pcseg_batch_stage_times pcseg_batch_profile pcseg_batch_device_labels pcseg_batch_device_normals
pcseg_regions_count pcseg_regions_inlier_offsets pcseg_regions_inlier_indices
pcseg_regions_label_ids pcseg_regions_seed_indices pcseg_regions_planes
pcseg_regions_centroids pcseg_regions_areas pcseg_regions_estimators
pcseg_regions_boundary_offsets pcseg_regions_boundary_indices
pcseg_regions_discontinuous_offsets pcseg_regions_discontinuous_indices
pcseg_regions_hull_offsets pcseg_regions_hull_points pcseg_regions_seed_points
pcseg_regions_free""".split()


class PcsegError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__("pcseg status %d: %s" % (status, msg))
        self.status = status


def lib():
    """Loads the CUDA library; raises if it has not been built (no fallback)."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libpcseg_b200.so is not built (run `make lib` or "
                               "__graft_entry__.build()); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        L.pcseg_status_string.restype = C.c_char_p
        L.pcseg_last_error.restype = C.c_char_p
        L.pcseg_last_error.argtypes = [C.c_void_p]
        L.pcseg_launch_count.restype = C.c_longlong
        L.pcseg_launch_count.argtypes = [C.c_void_p]
        L.pcseg_stream.restype = C.c_void_p
        L.pcseg_stream.argtypes = [C.c_void_p]
        L.pcseg_batch_stream.restype = C.c_void_p
        L.pcseg_batch_stream.argtypes = [C.c_void_p]
        L.pcseg_batch_wait.argtypes = [C.c_void_p]
        L.pcseg_destroy.argtypes = [C.c_void_p]
        L.pcseg_batch_destroy.argtypes = [C.c_void_p]
        L.pcseg_batch_device_labels.restype = C.c_void_p
        L.pcseg_batch_device_labels.argtypes = [C.c_void_p]
        L.pcseg_batch_device_normals.restype = C.c_void_p
        L.pcseg_batch_device_normals.argtypes = [C.c_void_p]
        for name, rt in [("inlier_offsets", I32P), ("inlier_indices", I32P),
                         ("label_ids", I32P), ("seed_indices", I32P),
                         ("planes", F32P), ("centroids", F32P), ("areas", F32P),
                         ("estimators", F32P), ("boundary_offsets", I32P),
                         ("boundary_indices", I32P),
                         ("discontinuous_offsets", I32P),
                         ("discontinuous_indices", I32P), ("hull_offsets", I32P),
                         ("hull_points", F32P), ("seed_points", F32P)]:
            f = getattr(L, "pcseg_regions_" + name)
            f.restype = rt
            f.argtypes = [C.c_void_p]
        L.pcseg_regions_count.argtypes = [C.c_void_p]
        L.pcseg_regions_free.argtypes = [C.c_void_p]
        _LIB = L
    return _LIB


def _collect(L, handle, kind):
    g = lambda n: getattr(L, "pcseg_regions_" + n)
    n = L.pcseg_regions_count(handle)

    def arr(name, count, dtype):
        if count == 0:
            return np.zeros(0, dtype=dtype)
        return np.ctypeslib.as_array(g(name)(handle), shape=(count,)).astype(dtype, copy=True)

    r = SimpleNamespace(n=n)
    r.inlier_offsets = arr("inlier_offsets", n + 1, np.int32)
    r.inlier_indices = arr("inlier_indices", int(r.inlier_offsets[-1]), np.int32)
    r.label_ids = arr("label_ids", n, np.int32)
    if kind == "planar":
        r.seed_indices = arr("seed_indices", n, np.int32)
        r.planes = arr("planes", 4 * n, np.float32).reshape(n, 4)
        r.centroids = arr("centroids", 3 * n, np.float32).reshape(n, 3)
        r.areas = arr("areas", n, np.float32)
        r.estimators = arr("estimators", 10 * n, np.float32).reshape(n, 10)
        r.boundary_offsets = arr("boundary_offsets", n + 1, np.int32)
        r.boundary_indices = arr("boundary_indices", int(r.boundary_offsets[-1]), np.int32)
        r.disc_offsets = arr("discontinuous_offsets", n + 1, np.int32)
        r.disc_indices = arr("discontinuous_indices", int(r.disc_offsets[-1]), np.int32)
        r.hull_offsets = arr("hull_offsets", n + 1, np.int32)
        r.hull_points = arr("hull_points", 3 * int(r.hull_offsets[-1]), np.float32).reshape(-1, 3)
    if kind == "meanshift":
        r.seed_points = arr("seed_points", 3 * n, np.float32).reshape(n, 3)
    r.inliers = [r.inlier_indices[r.inlier_offsets[i]:r.inlier_offsets[i + 1]] for i in range(n)]
    return r


def _addr(a):
    """Host numpy array -> pointer; int -> raw (device) address."""
    if isinstance(a, (int, np.integer)):
        return C.c_void_p(int(a))
    return C.c_void_p(a.ctypes.data)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def normals_params(**kw):
    p = NormalsParams()
    lib().pcseg_default_normals_params(C.byref(p))
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def planar_config(**kw):
    p = PlanarConfig()
    lib().pcseg_default_planar_config(C.byref(p))
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def cluster_config(**kw):
    p = ClusterConfig()
    lib().pcseg_default_cluster_config(C.byref(p))
    for k, v in kw.items():
        setattr(p, k, v)
    return p


class Context:
    def __init__(self, device=0):
        self.L = lib()
        self.h = C.c_void_p()
        st = self.L.pcseg_create(int(device), C.byref(self.h))
        if st != 0:
            raise PcsegError(st, "pcseg_create failed (no usable CUDA device; no CPU fallback)")

    def close(self):
        if self.h:
            self.L.pcseg_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, st):
        if st != 0:
            raise PcsegError(st, self.L.pcseg_status_string(st).decode() + " | " +
                             self.L.pcseg_last_error(self.h).decode())

    @property
    def launches(self):
        return int(self.L.pcseg_launch_count(self.h))

    @property
    def stream(self):
        return self.L.pcseg_stream(self.h)

    def prepare_unlabeled_labels(self, n):
        out = np.zeros(n, np.int32)
        self._check(self.L.pcseg_prepare_unlabeled_labels(self.h, _addr(out), n))
        return out

    def normals_organized(self, xyz, rows, cols, sensor_t=(0, 0, 0), params=None,
                          row_range=None, col_range=None, normals_init=None,
                          want_curvature=False):
        xyz = _f32(xyz).reshape(-1)
        p = params if isinstance(params, NormalsParams) else normals_params(**(params or {}))
        st = (C.c_double * 3)(*[float(v) for v in sensor_t])
        r0, r1 = row_range or (0, rows)
        c0, c1 = col_range or (0, cols)
        normals = (np.zeros(rows * cols * 3, np.float32) if normals_init is None
                   else _f32(normals_init).reshape(-1).copy())
        curv = np.zeros(rows * cols, np.float32) if want_curvature else None
        self._check(self.L.pcseg_normals_organized(
            self.h, _addr(xyz), rows, cols, st, C.byref(p), r0, r1, c0, c1,
            _addr(normals), _addr(curv) if want_curvature else None))
        normals = normals.reshape(rows * cols, 3)
        return (normals, curv) if want_curvature else normals

    def segment_planar(self, xyz, normals, rows, cols, seeds, labels, cfg=None,
                       id_offset=0, robot_q=(0, 0, 0, 1)):
        xyz = _f32(xyz).reshape(-1)
        normals = _f32(normals).reshape(-1)
        seeds = _i32(seeds)
        labels = _i32(labels).reshape(-1).copy()
        c = cfg if isinstance(cfg, PlanarConfig) else planar_config(**(cfg or {}))
        q = (C.c_float * 4)(*[float(v) for v in robot_q])
        h = C.c_void_p()
        self._check(self.L.pcseg_segment_planar(
            self.h, _addr(xyz), _addr(normals), rows, cols, q, _addr(seeds), seeds.size,
            C.byref(c), id_offset, _addr(labels), C.byref(h)))
        r = _collect(self.L, h, "planar")
        self.L.pcseg_regions_free(h)
        return labels, r

    def segment_cluster(self, xyz, rows, cols, seeds, labels, cfg=None, id_offset=0):
        xyz = _f32(xyz).reshape(-1)
        seeds = _i32(seeds)
        labels = _i32(labels).reshape(-1).copy()
        c = cfg if isinstance(cfg, ClusterConfig) else cluster_config(**(cfg or {}))
        h = C.c_void_p()
        self._check(self.L.pcseg_segment_cluster(
            self.h, _addr(xyz), rows, cols, _addr(seeds), seeds.size, C.byref(c),
            id_offset, _addr(labels), C.byref(h)))
        r = _collect(self.L, h, "cluster")
        self.L.pcseg_regions_free(h)
        return labels, r

    def mean_shift_modes(self, xyz, rows, cols, labels, iterations=5):
        xyz = _f32(xyz).reshape(-1)
        labels = _i32(labels).reshape(-1)
        out = np.zeros(rows * cols * 8, np.float32)
        self._check(self.L.pcseg_mean_shift_modes(self.h, _addr(xyz), rows, cols,
                                                  _addr(labels), iterations, _addr(out)))
        return out.reshape(-1, 8)

    def mean_shift_modes_range(self, xyz, rows, cols, labels, iterations, cell_begin, cell_end):
        """Mode records of the seed cells [cell_begin, cell_end) (one shard)."""
        xyz = _f32(xyz).reshape(-1)
        labels = _i32(labels).reshape(-1)
        out = np.zeros((cell_end - cell_begin) * 8, np.float32)
        self._check(self.L.pcseg_mean_shift_modes_range(self.h, _addr(xyz), rows, cols, _addr(labels),
                                                        iterations, cell_begin, cell_end, _addr(out)))
        return out.reshape(-1, 8)

    def mean_shift_from_modes(self, xyz, rows, cols, modes, labels, cfg=None, id_offset=0, sort_mode=0):
        xyz = _f32(xyz).reshape(-1)
        modes = _f32(modes).reshape(-1)
        labels = _i32(labels).reshape(-1).copy()
        c = cfg if isinstance(cfg, ClusterConfig) else cluster_config(**(cfg or {}))
        h = C.c_void_p()
        self._check(self.L.pcseg_mean_shift_from_modes(self.h, _addr(xyz), rows, cols, _addr(modes), C.byref(c),
                                                       id_offset, sort_mode, _addr(labels), C.byref(h)))
        r = _collect(self.L, h, "meanshift")
        self.L.pcseg_regions_free(h)
        return labels, r

    def mean_shift(self, xyz, rows, cols, labels, cfg=None, iterations=5, id_offset=0, sort_mode=0):
        """SlidingMeanShift (ps/mean_shift_segmentation.h:207-330): labels out, regions."""
        xyz = _f32(xyz).reshape(-1)
        labels = _i32(labels).reshape(-1).copy()
        c = cfg if isinstance(cfg, ClusterConfig) else cluster_config(**(cfg or {}))
        h = C.c_void_p()
        self._check(self.L.pcseg_mean_shift(self.h, _addr(xyz), rows, cols, C.byref(c), iterations,
                                            id_offset, sort_mode, _addr(labels), C.byref(h)))
        r = _collect(self.L, h, "meanshift")
        self.L.pcseg_regions_free(h)
        return labels, r

    def seeds_plane_support(self, xyz, normals, rows, cols, neighborhood_size=9,
                            max_plane_distance=0.05, min_support=12):
        xyz = _f32(xyz).reshape(-1)
        normals = _f32(normals).reshape(-1)
        out = np.zeros(rows * cols, np.int32)
        n = C.c_int(0)
        self._check(self.L.pcseg_seeds_plane_support(
            self.h, _addr(xyz), _addr(normals), rows, cols, neighborhood_size,
            C.c_float(max_plane_distance), min_support, _addr(out), out.size, C.byref(n)))
        return out[:n.value].copy()


class Batch:
    """Persistent workspace for n_frames x rows x cols (pcseg_batch_*)."""

    def __init__(self, ctx, n_frames, rows, cols):
        self.ctx, self.n_frames, self.rows, self.cols = ctx, n_frames, rows, cols
        self.h = C.c_void_p()
        ctx._check(ctx.L.pcseg_batch_create(ctx.h, n_frames, rows, cols, C.byref(self.h)))

    def close(self):
        if self.h:
            self.ctx.L.pcseg_batch_destroy(self.h)
            self.h = C.c_void_p()

    def run(self, xyz, params, normals_out=None, labels_out=None):
        """xyz / outputs: host numpy arrays or raw device addresses (int)."""
        self.ctx._check(self.ctx.L.pcseg_batch_run(
            self.h, _addr(xyz), C.byref(params),
            _addr(normals_out) if normals_out is not None else None,
            _addr(labels_out) if labels_out is not None else None))

    def submit(self, xyz, params, normals_out=None, labels_out=None):
        """Asynchronous run (pcseg_batch_submit); the buffers must stay alive until wait()."""
        self._keep = (xyz, params, normals_out, labels_out)
        self.ctx._check(self.ctx.L.pcseg_batch_submit(
            self.h, _addr(xyz), C.byref(params),
            _addr(normals_out) if normals_out is not None else None,
            _addr(labels_out) if labels_out is not None else None))

    def wait(self):
        self.ctx._check(self.ctx.L.pcseg_batch_wait(self.h))
        self._keep = None

    @property
    def stream(self):
        return self.ctx.L.pcseg_batch_stream(self.h)

    def frame_stats(self):
        out = np.zeros((self.n_frames, 8), np.int32)
        self.ctx._check(self.ctx.L.pcseg_batch_frame_stats(self.h, _addr(out)))
        return out

    def stage_times(self):
        out = np.zeros(4, np.float32)
        self.ctx._check(self.ctx.L.pcseg_batch_stage_times(self.h, _addr(out)))
        return dict(zip(("normals", "seeds", "planar", "cluster"), [float(v) for v in out]))

    def profile(self):
        out = np.zeros((self.n_frames, 16), np.int64)
        self.ctx._check(self.ctx.L.pcseg_batch_profile(self.h, _addr(out)))
        return out

    def regions(self, frame, kind):
        h = C.c_void_p()
        self.ctx._check(self.ctx.L.pcseg_batch_regions(self.h, frame, 0 if kind == "planar" else 1, C.byref(h)))
        r = _collect(self.ctx.L, h, kind)
        self.ctx.L.pcseg_regions_free(h)
        return r

    @property
    def device_labels(self):
        return self.ctx.L.pcseg_batch_device_labels(self.h)

    @property
    def device_normals(self):
        return self.ctx.L.pcseg_batch_device_normals(self.h)


def pipeline_params(normals=None, planar=None, cluster=None, sensor_t=(0, 0, 0),
                    robot_q=(0, 0, 0, 1), seed_mode=SEEDS_ALL_DESCENDING,
                    seeds=None, run_cluster=True):
    p = PipelineParams()
    p.normals = normals or normals_params()
    p.planar = planar or planar_config()
    p.cluster = cluster or cluster_config()
    for i in range(3):
        p.sensor_translation[i] = float(sensor_t[i])
    for i in range(4):
        p.robot_q_xyzw[i] = float(robot_q[i])
    p.planar_seed_mode = seed_mode
    if seeds is not None:
        p._keep = _i32(seeds)
        p.planar_seeds = p._keep.ctypes.data_as(I32P)
        p.n_planar_seeds = p._keep.size
    p.run_cluster_stage = int(run_cluster)
    return p
