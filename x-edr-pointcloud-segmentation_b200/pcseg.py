"""ctypes binding of libpcseg_b200.so (the C ABI in include/pcseg_b200.h).

Thin host-side mirror used by tests/, bench.py and __graft_entry__.py.  There
is NO CPU fallback: if the CUDA library is missing or no GPU is present,
loading / Context() raises.
"""
import ctypes as C
import os
from types import SimpleNamespace

import numpy as np

# Several workspaces (each with its own streams) overlap only if their streams map to distinct
# hardware queues; the default is 8.  Read by the driver when the CUDA context is created.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

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


class PlanarRecord(C.Structure):
    _fields_ = [("frame", C.c_int), ("label_id", C.c_int), ("seed_index", C.c_int),
                ("inlier_begin", C.c_int), ("inlier_count", C.c_int),
                ("boundary_begin", C.c_int), ("boundary_count", C.c_int),
                ("discontinuous_begin", C.c_int), ("discontinuous_count", C.c_int),
                ("hull_begin", C.c_int), ("hull_count", C.c_int),
                ("plane", C.c_float * 4), ("centroid", C.c_float * 3), ("area", C.c_float),
                ("estimator", C.c_float * 10)]


class ClusterRecord(C.Structure):
    _fields_ = [("frame", C.c_int), ("label_id", C.c_int), ("inlier_begin", C.c_int), ("inlier_count", C.c_int),
                ("centroid", C.c_float * 3), ("bbox_min", C.c_float * 3), ("bbox_max", C.c_float * 3)]


PLANAR_RECORD_DTYPE = np.dtype([
    ("frame", "<i4"), ("label_id", "<i4"), ("seed_index", "<i4"), ("inlier_begin", "<i4"), ("inlier_count", "<i4"),
    ("boundary_begin", "<i4"), ("boundary_count", "<i4"), ("discontinuous_begin", "<i4"),
    ("discontinuous_count", "<i4"), ("hull_begin", "<i4"), ("hull_count", "<i4"),
    ("plane", "<f4", 4), ("centroid", "<f4", 3), ("area", "<f4"), ("estimator", "<f4", 10)])
CLUSTER_RECORD_DTYPE = np.dtype([
    ("frame", "<i4"), ("label_id", "<i4"), ("inlier_begin", "<i4"), ("inlier_count", "<i4"),
    ("centroid", "<f4", 3), ("bbox_min", "<f4", 3), ("bbox_max", "<f4", 3)])
assert PLANAR_RECORD_DTYPE.itemsize == C.sizeof(PlanarRecord) and CLUSTER_RECORD_DTYPE.itemsize == C.sizeof(ClusterRecord)


class BatchResultsStruct(C.Structure):
    _fields_ = [("want_inlier_xyz", C.c_int), ("want_boundary_xyz", C.c_int), ("want_hull", C.c_int),
                ("want_cluster_stats", C.c_int),
                ("frame_stats", C.c_void_p), ("frame_offsets", C.c_void_p),
                ("planar", C.c_void_p), ("planar_cap", C.c_int),
                ("planar_inlier_indices", C.c_void_p), ("planar_inlier_cap", C.c_int),
                ("planar_inlier_xyz", C.c_void_p),
                ("boundary_indices", C.c_void_p), ("boundary_cap", C.c_int),
                ("boundary_xyz", C.c_void_p), ("boundary_discontinuous", C.c_void_p),
                ("discontinuous_indices", C.c_void_p), ("discontinuous_positions", C.c_void_p),
                ("hull_points", C.c_void_p),
                ("cluster", C.c_void_p), ("cluster_cap", C.c_int),
                ("cluster_inlier_indices", C.c_void_p), ("cluster_inlier_cap", C.c_int),
                ("cluster_inlier_xyz", C.c_void_p),
                ("n_planar", C.c_int), ("n_planar_inliers", C.c_int), ("n_boundary", C.c_int),
                ("n_cluster", C.c_int), ("n_cluster_inliers", C.c_int), ("error_flags", C.c_int)]


SEEDS_ALL_DESCENDING, SEEDS_PLANE_SUPPORT, SEEDS_EXPLICIT = 0, 1, 2

EXPORTS = """pcseg_default_normals_params pcseg_default_planar_config
pcseg_default_cluster_config pcseg_create pcseg_destroy pcseg_status_string
pcseg_last_error pcseg_launch_count pcseg_stream pcseg_prepare_unlabeled_labels
pcseg_normals_organized pcseg_segment_planar pcseg_segment_cluster
pcseg_mean_shift_modes pcseg_mean_shift_modes_range pcseg_mean_shift_from_modes
pcseg_mean_shift pcseg_mean_shift_sort_order pcseg_seeds_plane_support pcseg_seeds_average_normals pcseg_seeds_average_normals_noalloc pcseg_seeds_last_planar_regions
pcseg_fit_planes pcseg_plane_estimator_replay pcseg_plane_estimators_transform pcseg_plane_estimators_merge
pcseg_batch_submit_regions pcseg_transform_in_place pcseg_gather_points pcseg_cluster_unorganized pcseg_regions_bboxes
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
                         ("hull_points", F32P), ("seed_points", F32P), ("bboxes", F32P)]:
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
    if kind == "unorganized":
        r.centroids = arr("centroids", 3 * n, np.float32).reshape(n, 3)
        r.bboxes = arr("bboxes", 6 * n, np.float32).reshape(n, 6)
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

    def cluster_unorganized(self, xyz, labels, radius=0.5, min_region_inliers=10, id_offset=0):
        """Fixed-radius Euclidean clustering of an unorganized cloud through a voxel-hash grid
        (pcseg_cluster_unorganized; extension): labels out, regions with
        inlier lists, centroids and extents."""
        xyz = _f32(xyz).reshape(-1)
        labels = _i32(labels).reshape(-1).copy()

        class Cfg(C.Structure):
            _fields_ = [("radius", C.c_float), ("min_region_inliers", C.c_int)]
        cfg = Cfg(radius, min_region_inliers)
        h = C.c_void_p()
        self._check(self.L.pcseg_cluster_unorganized(self.h, _addr(xyz), labels.size, C.byref(cfg), int(id_offset),
                                                     _addr(labels), C.byref(h)))
        r = _collect(self.L, h, "unorganized")
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

    def mean_shift_sort_order(self, intensities, sort_mode=0):
        """Positions of the valid seeds (intensity >= 0) in the order of the reference's sort."""
        v = _f32(intensities).reshape(-1)
        out = np.zeros(max(v.size, 1), np.int32)
        n = C.c_int(0)
        self._check(self.L.pcseg_mean_shift_sort_order(self.h, _addr(v), v.size, int(sort_mode), _addr(out), C.byref(n)))
        return out[:n.value].copy()

    def seeds_average_normals(self, xyz, normals, rows, cols, neighborhood_size=5,
                              min_num_valid_normals=8, min_avg_normal_length=0.9999):
        """FindSeedPointsFromAverageNormals (ps/segmentation.h:137-184)."""
        xyz = _f32(xyz).reshape(-1)
        normals = _f32(normals).reshape(-1)
        out = np.zeros(rows * cols, np.int32)
        n = C.c_int(0)
        self._check(self.L.pcseg_seeds_average_normals(
            self.h, _addr(xyz), _addr(normals), rows, cols, neighborhood_size, min_num_valid_normals,
            C.c_float(min_avg_normal_length), _addr(out), out.size, C.byref(n)))
        return out[:n.value].copy()

    def seeds_average_normals_noalloc(self, normals_dev, rows, cols, seeds_out_dev, cap, n_out_dev,
                                      neighborhood_size=5, min_num_valid_normals=8,
                                      min_avg_normal_length=0.9999):
        """pcseg_seeds_average_normals_noalloc: the same seeds with the scratch in a static device
        array (no allocation inside the call).  normals_dev / seeds_out_dev / n_out_dev are DEVICE
        addresses (ints) owned by the caller; returns after the context's stream has drained."""
        self._check(self.L.pcseg_seeds_average_normals_noalloc(
            self.h, C.c_void_p(normals_dev), rows, cols, neighborhood_size, min_num_valid_normals,
            C.c_float(min_avg_normal_length), C.c_void_p(seeds_out_dev), cap, C.c_void_p(n_out_dev)))

    def seeds_last_planar_regions(self, xyz, normals, prev_inlier_counts, prev_centroids, prev_normals,
                                  q_xyzw=(0, 0, 0, 1), translation=(0, 0, 0), max_distance=0.5,
                                  max_normal_difference_angle=0.2, seeds=None):
        """FindSeedPointsFromLastPlanarRegions (ps/planar_region.h:478-519); appends to `seeds`."""
        xyz = _f32(xyz).reshape(-1)
        normals = _f32(normals).reshape(-1)
        cnt = _i32(prev_inlier_counts)
        cen = _f32(prev_centroids).reshape(-1)
        nrm = _f32(prev_normals).reshape(-1)
        have = _i32(seeds if seeds is not None else [])
        out = np.zeros(have.size + max(cnt.size, 1), np.int32)
        out[:have.size] = have
        n = C.c_int(have.size)
        q = (C.c_float * 4)(*[float(v) for v in q_xyzw])
        t = (C.c_float * 3)(*[float(v) for v in translation])
        self._check(self.L.pcseg_seeds_last_planar_regions(
            self.h, _addr(xyz), _addr(normals), xyz.size // 3, cnt.size, _addr(cnt), _addr(cen), _addr(nrm), q, t,
            C.c_float(max_distance), C.c_float(max_normal_difference_angle), _addr(out), out.size, C.byref(n)))
        return out[:n.value].copy()

    # ---- PlaneEstimator (ps/plane_estimator.h:37-120) ----
    def fit_planes(self, xyz, offsets, indices, weights=None, hints=None, want_state=False):
        """Plane fits over CSR index sets; returns query rows [n, 10] (and states [n, 10])."""
        xyz = _f32(xyz).reshape(-1)
        offsets = _i32(offsets)
        indices = _i32(indices)
        n = offsets.size - 1
        w = _f32(weights) if weights is not None else None
        h = _f32(hints).reshape(-1) if hints is not None else None
        q = np.zeros((max(n, 0), 10), np.float32)
        st = np.zeros((max(n, 0), 10), np.float32) if want_state else None
        self._check(self.L.pcseg_fit_planes(
            self.h, _addr(xyz), xyz.size // 3, _addr(offsets), _addr(indices) if indices.size else None, n,
            _addr(w) if w is not None else None, _addr(h) if h is not None else None, _addr(q),
            _addr(st) if want_state else None))
        return (q, st) if want_state else q

    def plane_estimator_replay(self, ops, poses=()):
        """Same op list as oracle.plane_estimator_replay2; returns query rows [n_queries, 10]."""
        codes = _i32([o[0] for o in ops])
        args = _f32([[o[1], o[2], o[3], o[4]] for o in ops]).reshape(-1)
        pz = np.ascontiguousarray(np.asarray(poses, np.float64).reshape(-1))
        cap = max(1, sum(1 for o in ops if o[0] == 3))
        out = np.zeros((cap, 10), np.float32)
        n = C.c_int(0)
        self._check(self.L.pcseg_plane_estimator_replay(
            self.h, _addr(codes), _addr(args), len(ops), _addr(pz) if pz.size else None, pz.size // 7, _addr(out),
            cap, C.byref(n)))
        return out[:n.value].copy()

    def plane_estimators_transform(self, states, poses):
        """Transform of [n, 10] accumulator states (returns a copy); poses [n, 7] or one (7,)."""
        s = _f32(states).reshape(-1, 10).copy()
        pz = np.ascontiguousarray(np.asarray(poses, np.float64).reshape(-1))
        self._check(self.L.pcseg_plane_estimators_transform(self.h, _addr(s), s.shape[0], _addr(pz), int(pz.size > 7)))
        return s

    def plane_estimators_merge(self, dst, src, poses=None):
        d = _f32(dst).reshape(-1, 10).copy()
        s = _f32(src).reshape(-1, 10)
        pz = np.ascontiguousarray(np.asarray(poses, np.float64).reshape(-1)) if poses is not None else None
        self._check(self.L.pcseg_plane_estimators_merge(
            self.h, _addr(d), _addr(s), d.shape[0], _addr(pz) if pz is not None else None,
            int(pz is not None and pz.size > 7)))
        return d

    def transform_in_place(self, points=None, normals=None, q_xyzw=(0, 0, 0, 1), translation=(0, 0, 0)):
        """MultichannelCloud::TransformInPlace (ps/multichannel_cloud.cc:53-68); arrays are modified in place."""
        n = (points if points is not None else normals).reshape(-1, 3).shape[0]
        q = (C.c_double * 4)(*[float(v) for v in q_xyzw])
        t = (C.c_double * 3)(*[float(v) for v in translation])
        self._check(self.L.pcseg_transform_in_place(
            self.h, _addr(points) if points is not None else None,
            _addr(normals) if normals is not None else None, n, q, t))

    def gather_points(self, xyz, indices, discontinuous=None):
        """Create*DetectedObjectProto gathers (ps/detected_objects.cc:36-46, .h:62-74)."""
        xyz = _f32(xyz).reshape(-1)
        indices = _i32(indices)
        out = np.zeros((indices.size, 3), np.float32)
        disc = _i32(discontinuous) if discontinuous is not None else np.zeros(0, np.int32)
        pos = np.zeros(max(indices.size, 1), np.int32)
        npos = C.c_int(0)
        self._check(self.L.pcseg_gather_points(
            self.h, _addr(xyz), xyz.size // 3, _addr(indices), indices.size, _addr(out),
            _addr(disc) if disc.size else None, disc.size, _addr(pos), C.byref(npos)))
        return out, pos[:npos.value].copy()


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
        out = np.zeros(5, np.float32)
        self.ctx._check(self.ctx.L.pcseg_batch_stage_times(self.h, _addr(out)))
        return dict(zip(("normals", "seeds", "planar", "cluster", "pack"), [float(v) for v in out]))

    def submit_regions(self, xyz, params, results, normals_out=None, labels_out=None):
        """Asynchronous run that also returns every frame's regions (pcseg_batch_submit_regions);
        `results` is a BatchResults whose buffers must stay alive until wait()."""
        self._keep = (xyz, params, normals_out, labels_out, results)
        self.ctx._check(self.ctx.L.pcseg_batch_submit_regions(
            self.h, _addr(xyz), C.byref(params),
            _addr(normals_out) if normals_out is not None else None,
            _addr(labels_out) if labels_out is not None else None, C.byref(results.s)))

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


class BatchResults:
    """Caller-side buffers of pcseg_batch_results for a workspace of n_frames x n_points.
    alloc(shape, dtype) -> array-like with .ctypes / numpy view; pass bench.py's pinned allocator for
    overlap, default numpy.  Capacities: every point can be an inlier; boundary entries are bounded
    by boundary_cap (default n_frames * 8192)."""

    def __init__(self, n_frames, n_points, inlier_xyz=False, boundary_xyz=False, hull=True, cluster_stats=True,
                 boundary_cap=None, region_cap=None, alloc=None, share=None):
        base = alloc or (lambda shape, dtype: np.zeros(shape, dtype))
        names = iter(("frame_stats", "frame_offsets", "planar", "planar_inlier_indices", "planar_inlier_xyz",
                      "boundary_indices", "boundary_xyz", "boundary_discontinuous", "discontinuous_indices",
                      "discontinuous_positions", "hull_points", "cluster", "cluster_inlier_indices",
                      "cluster_inlier_xyz"))
        skipped = {"planar_inlier_xyz": not inlier_xyz, "boundary_xyz": not boundary_xyz, "hull_points": not hull,
                   "cluster_inlier_xyz": not inlier_xyz}

        def alloc(shape, dtype):
            """Arrays are requested in the order of `names`; one that `share` (another BatchResults of
            the same geometry) already holds is reused instead of allocated."""
            name = next(names)
            while skipped.get(name):
                name = next(names)
            have = getattr(share, name, None) if share is not None else None
            return have if have is not None else base(shape, dtype)

        F, N = n_frames, n_points
        bc = boundary_cap or F * 8192
        rc = region_cap or F * 256
        self.F = F
        self.frame_stats = alloc((F, 8), np.int32)
        self.frame_offsets = alloc((5, F + 1), np.int32)
        self.planar = alloc((rc,), PLANAR_RECORD_DTYPE)
        self.planar_inlier_indices = alloc((F * N,), np.int32)
        self.planar_inlier_xyz = alloc((F * N, 3), np.float32) if inlier_xyz else None
        self.boundary_indices = alloc((bc,), np.int32)
        self.boundary_xyz = alloc((bc, 3), np.float32) if boundary_xyz else None
        self.boundary_discontinuous = alloc((bc,), np.uint8)
        self.discontinuous_indices = alloc((bc,), np.int32)
        self.discontinuous_positions = alloc((bc,), np.int32)
        self.hull_points = alloc((bc, 3), np.float32) if hull else None
        self.cluster = alloc((rc,), CLUSTER_RECORD_DTYPE)
        self.cluster_inlier_indices = alloc((F * N,), np.int32)
        self.cluster_inlier_xyz = alloc((F * N, 3), np.float32) if inlier_xyz else None
        s = BatchResultsStruct()
        s.want_inlier_xyz, s.want_boundary_xyz, s.want_hull = int(inlier_xyz), int(boundary_xyz), int(hull)
        s.want_cluster_stats = int(cluster_stats)
        ptr = lambda a: (a.ctypes.data if a is not None else None)
        s.frame_stats, s.frame_offsets = ptr(self.frame_stats), ptr(self.frame_offsets)
        s.planar, s.planar_cap = ptr(self.planar), rc
        s.planar_inlier_indices, s.planar_inlier_cap = ptr(self.planar_inlier_indices), F * N
        s.planar_inlier_xyz = ptr(self.planar_inlier_xyz)
        s.boundary_indices, s.boundary_cap = ptr(self.boundary_indices), bc
        s.boundary_xyz, s.boundary_discontinuous = ptr(self.boundary_xyz), ptr(self.boundary_discontinuous)
        s.discontinuous_indices, s.discontinuous_positions = ptr(self.discontinuous_indices), ptr(self.discontinuous_positions)
        s.hull_points = ptr(self.hull_points)
        s.cluster, s.cluster_cap = ptr(self.cluster), rc
        s.cluster_inlier_indices, s.cluster_inlier_cap = ptr(self.cluster_inlier_indices), F * N
        s.cluster_inlier_xyz = ptr(self.cluster_inlier_xyz)
        self.s = s

    @property
    def d2h_bytes(self):
        """Bytes the last completed run copied into these buffers."""
        s = self.s
        n = 4 * (5 * (self.F + 1) + 8 * self.F)
        n += s.n_planar * PLANAR_RECORD_DTYPE.itemsize + s.n_cluster * CLUSTER_RECORD_DTYPE.itemsize
        n += 4 * s.n_planar_inliers + 4 * s.n_cluster_inliers
        if s.want_inlier_xyz:
            n += 12 * (s.n_planar_inliers + s.n_cluster_inliers)
        n += s.n_boundary * (4 + 1 + 4 + 4 + (12 if s.want_boundary_xyz else 0) + (12 if s.want_hull else 0))
        return n

    def frame_regions(self, f, kind):
        """Regions of frame f as the same namespace _collect() builds (views, frame-local CSR)."""
        s = self.s
        off = np.asarray(self.frame_offsets)
        r = SimpleNamespace()
        if kind == "planar":
            recs = np.asarray(self.planar)[off[0, f]:off[0, f + 1]]
            r.n = len(recs)
            ii = np.asarray(self.planar_inlier_indices)
            r.inliers = [ii[m["inlier_begin"]:m["inlier_begin"] + m["inlier_count"]] for m in recs]
            r.label_ids = recs["label_id"].copy()
            r.seed_indices = recs["seed_index"].copy()
            r.planes, r.centroids, r.areas, r.estimators = recs["plane"], recs["centroid"], recs["area"], recs["estimator"]
            bi = np.asarray(self.boundary_indices)
            r.boundaries = [bi[m["boundary_begin"]:m["boundary_begin"] + m["boundary_count"]] for m in recs]
            di, dp = np.asarray(self.discontinuous_indices), np.asarray(self.discontinuous_positions)
            r.disc = [di[m["discontinuous_begin"]:m["discontinuous_begin"] + m["discontinuous_count"]] for m in recs]
            r.disc_positions = [dp[m["discontinuous_begin"]:m["discontinuous_begin"] + m["discontinuous_count"]] for m in recs]
            if self.hull_points is not None:
                hp = np.asarray(self.hull_points)
                r.hulls = [hp[m["hull_begin"]:m["hull_begin"] + m["hull_count"]] for m in recs]
            if self.planar_inlier_xyz is not None:
                px = np.asarray(self.planar_inlier_xyz)
                r.inlier_xyz = [px[m["inlier_begin"]:m["inlier_begin"] + m["inlier_count"]] for m in recs]
            if self.boundary_xyz is not None:
                bx = np.asarray(self.boundary_xyz)
                r.boundary_xyz = [bx[m["boundary_begin"]:m["boundary_begin"] + m["boundary_count"]] for m in recs]
        else:
            recs = np.asarray(self.cluster)[off[3, f]:off[3, f + 1]]
            r.n = len(recs)
            ii = np.asarray(self.cluster_inlier_indices)
            r.inliers = [ii[m["inlier_begin"]:m["inlier_begin"] + m["inlier_count"]] for m in recs]
            r.label_ids = recs["label_id"].copy()
            r.centroids, r.bbox_min, r.bbox_max = recs["centroid"], recs["bbox_min"], recs["bbox_max"]
            if self.cluster_inlier_xyz is not None:
                cx = np.asarray(self.cluster_inlier_xyz)
                r.inlier_xyz = [cx[m["inlier_begin"]:m["inlier_begin"] + m["inlier_count"]] for m in recs]
        r.records = recs
        return r


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
