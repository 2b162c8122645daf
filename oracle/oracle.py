"""ctypes binding of the CPU ORACLE (oracle/liboracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product never imports it.
"""
import ctypes as C
import os
import subprocess
from types import SimpleNamespace

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
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


def default_normals_params(**kw):
    """ps/algorithms.h:313-322"""
    d = dict(min_neighbor_distance=0.1, max_neighbor_distance=1.0,
             include_diagonal_neighbors=1, min_num_support_neighbors=4)
    d.update(kw)
    return d


def default_planar_config(**kw):
    """ps/planar_region.h:93-121"""
    d = dict(max_plane_distance=0.05, min_region_area=0.05, min_region_inliers=5,
             plane_model_reestimation_period=30, discontinuity_min_range=1.2,
             discontinuity_max_range=4.0, discontinuity_normal_angle_diff=5.0,
             discontinuity_z_diff=0.05, discontinuity_z_ratio=0.7)
    d.update(kw)
    return d


def default_cluster_config(**kw):
    """ps/cluster_region.h:53-63"""
    d = dict(min_region_inliers=7, squared_distance_threshold=1.0,
             half_search_window=1)
    d.update(kw)
    return d


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "pcseg_oracle.cc")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"],
                              stdout=subprocess.DEVNULL)
    return so


def _bind(so):
    L = C.CDLL(so)
    L.oracle_polygon_area.restype = C.c_float
    for name, rt in [("inlier_offsets", I32P), ("inlier_indices", I32P),
                     ("label_ids", I32P), ("seed_indices", I32P),
                     ("planes", F32P), ("centroids", F32P), ("areas", F32P),
                     ("estimators", F32P), ("boundary_offsets", I32P),
                     ("boundary_indices", I32P),
                     ("discontinuous_offsets", I32P),
                     ("discontinuous_indices", I32P), ("hull_offsets", I32P),
                     ("hull_points", F32P), ("seed_points", F32P)]:
        f = getattr(L, "oracle_regions_" + name)
        f.restype = rt
        f.argtypes = [C.c_void_p]
    L.oracle_regions_count.argtypes = [C.c_void_p]
    L.oracle_regions_free.argtypes = [C.c_void_p]
    return L


_REF = None       # oracle/_ref/libref.so: the reference's own sources against shims
_USE_REF = False


def ref_path():
    return os.path.join(_HERE, "_ref", "libref.so")


def ref_available():
    """True when oracle/_ref/libref.so exists (built here from /root/reference by
    `make -C oracle ref`; it travels to the GPU box as a prebuilt file)."""
    return os.path.exists(ref_path())


class reference:
    """Context manager: inside it every function of this module runs the
    REFERENCE's own code (oracle/_ref) instead of the restatement."""

    def __enter__(self):
        global _USE_REF
        self.prev = _USE_REF
        _USE_REF = True
        return self

    def __exit__(self, *a):
        global _USE_REF
        _USE_REF = self.prev


def lib():
    global _LIB, _REF
    if _USE_REF:
        if _REF is None:
            _REF = _bind(ref_path())
        return _REF
    if _LIB is None:
        so = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(so):
            build()
        _LIB = _bind(so)
    return _LIB


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _ptr(a, t):
    return a.ctypes.data_as(t)


def collect_regions(L, prefix, handle, kind):
    """Copies a regions handle into numpy arrays (shared by oracle & product)."""
    g = lambda n: getattr(L, prefix + "regions_" + n)
    n = g("count")(handle)

    def arr(name, count, dtype):
        if count == 0:
            return np.zeros(0, dtype=dtype)
        p = g(name)(handle)
        return np.ctypeslib.as_array(p, shape=(count,)).astype(dtype, copy=True)

    r = SimpleNamespace()
    r.n = n
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


def set_native_libm(on):
    lib().oracle_set_native_libm(int(on))


def normals_organized(xyz, rows, cols, sensor_t=(0, 0, 0), params=None,
                      row_range=None, col_range=None, normals_init=None,
                      want_curvature=False):
    L = lib()
    xyz = _f32(xyz).reshape(-1)
    p = NormalsParams(**(params or default_normals_params()))
    st = (C.c_double * 3)(*[float(v) for v in sensor_t])
    r0, r1 = row_range or (0, rows)
    c0, c1 = col_range or (0, cols)
    if normals_init is None:
        normals = np.zeros(rows * cols * 3, np.float32)
    else:
        normals = _f32(normals_init).reshape(-1).copy()
    curv = np.zeros(rows * cols, np.float32) if want_curvature else None
    L.oracle_normals_organized(_ptr(xyz, F32P), rows, cols, st, C.byref(p),
                               r0, r1, c0, c1, _ptr(normals, F32P),
                               _ptr(curv, F32P) if want_curvature else None)
    normals = normals.reshape(rows * cols, 3)
    return (normals, curv) if want_curvature else normals


def find_normal_support_neighbors(xyz, rows, cols, col, row, min_d, max_d, diag=True):
    L = lib()
    xyz = _f32(xyz).reshape(-1)
    out = np.zeros(9, np.int32)
    n = L.oracle_find_normal_support_neighbors(_ptr(xyz, F32P), rows, cols, col,
                                               row, C.c_float(min_d),
                                               C.c_float(max_d), int(diag),
                                               _ptr(out, I32P))
    return out[:n].copy()


def plane_estimator_replay(ops):
    """ops: list of (op, x, y, z, w). Returns array [n_queries, 10]."""
    L = lib()
    codes = _i32([o[0] for o in ops])
    args = _f32([[o[1], o[2], o[3], o[4]] for o in ops]).reshape(-1)
    nq = sum(1 for o in ops if o[0] == 3)
    out = np.zeros(max(nq, 1) * 10, np.float32)
    got = L.oracle_plane_estimator_replay(_ptr(codes, I32P), _ptr(args, F32P),
                                          len(ops), _ptr(out, F32P), out.size)
    assert got == nq, got
    return out.reshape(-1, 10)[:nq]


def plane_estimator_replay2(ops, poses=()):
    """Extended op list (see pcseg_oracle.h): 5 select slot, 6 Merge(slot), 7 Transform(pose),
    8 Merge(slot, pose).  poses: list of (qx, qy, qz, qw, tx, ty, tz) doubles."""
    L = lib()
    codes = _i32([o[0] for o in ops])
    args = _f32([[o[1], o[2], o[3], o[4]] for o in ops]).reshape(-1)
    pz = np.ascontiguousarray(np.asarray(poses, np.float64).reshape(-1))
    nq = sum(1 for o in ops if o[0] == 3)
    out = np.zeros(max(nq, 1) * 10, np.float32)
    got = L.oracle_plane_estimator_replay2(_ptr(codes, I32P), _ptr(args, F32P), len(ops),
                                           pz.ctypes.data_as(C.POINTER(C.c_double)), len(pz) // 7,
                                           _ptr(out, F32P), out.size)
    if got < 0:
        raise ValueError("oracle_plane_estimator_replay2 status %d" % got)
    assert got == nq, got
    return out.reshape(-1, 10)[:nq]


def fit_planes_csr(xyz, offsets, indices, hints=None, weights=None):
    """PlaneEstimator over CSR index sets (ps/plane_estimator.h:97-103 AddPoints(points, indices) per
    region, after SetNormalOrientation(hint) when hints are given; weighted AddPoint when weights are).
    Returns [n, 10] query rows {valid, curvature, centroid xyz, normal xyz, offset, weight}; a region
    with no points gives the state right after Clear()."""
    xyz = _f32(xyz).reshape(-1, 3)
    offsets = np.asarray(offsets)
    rows = []
    for r in range(len(offsets) - 1):
        ops = []
        if hints is not None:
            ops.append((2, *[float(v) for v in hints[r]], 0.0))
        for k in range(int(offsets[r]), int(offsets[r + 1])):
            pt = xyz[int(indices[k])]
            if weights is None:
                ops.append((0, float(pt[0]), float(pt[1]), float(pt[2]), 0.0))
            else:
                ops.append((1, float(pt[0]), float(pt[1]), float(pt[2]), float(weights[k])))
        ops.append((3, 0, 0, 0, 0))
        rows.append(plane_estimator_replay(ops)[0])
    return np.asarray(rows, np.float32).reshape(-1, 10)


def seeds_last_planar_regions(xyz, normals, prev_inlier_counts, prev_centroids, prev_normals,
                              q_xyzw=(0, 0, 0, 1), translation=(0, 0, 0), max_distance=0.5,
                              max_normal_difference_angle=0.2):
    """ps/planar_region.h:478-519; returns the appended seed indices."""
    L = lib()
    xyz = _f32(xyz).reshape(-1)
    normals = _f32(normals).reshape(-1)
    cnt = _i32(prev_inlier_counts)
    cen = _f32(prev_centroids).reshape(-1)
    nrm = _f32(prev_normals).reshape(-1)
    q = (C.c_float * 4)(*[float(v) for v in q_xyzw])
    t = (C.c_float * 3)(*[float(v) for v in translation])
    out = np.zeros(max(len(cnt), 1), np.int32)
    n = L.oracle_seeds_last_planar_regions(_ptr(xyz, F32P), _ptr(normals, F32P), xyz.size // 3, len(cnt),
                                           _ptr(cnt, I32P), _ptr(cen, F32P), _ptr(nrm, F32P), q, t,
                                           C.c_float(max_distance), C.c_float(max_normal_difference_angle),
                                           _ptr(out, I32P))
    return out[:n].copy()


def sort_order(intensities, sort_mode=0):
    """Permutation of the reference's std::sort by intensity (ps/mean_shift_segmentation.h:276-280)."""
    L = port_lib()
    v = np.ascontiguousarray(intensities, np.float64)
    out = np.zeros(v.size, np.int32)
    L.oracle_sort_order(v.ctypes.data_as(C.POINTER(C.c_double)), v.size, int(sort_mode), _ptr(out, I32P))
    return out


def eig3(a6):
    L = lib()
    a = _f32(a6)
    ev = np.zeros(3, np.float32)
    evec = np.zeros(9, np.float32)
    L.oracle_eig3(_ptr(a, F32P), _ptr(ev, F32P), _ptr(evec, F32P))
    return ev, evec.reshape(3, 3).T.copy()  # [row, col]


def polygon_area(xyz, indices=None):
    L = lib()
    xyz = _f32(xyz).reshape(-1)
    n_points = xyz.size // 3
    cn = np.zeros(3, np.float32)
    if indices is None:
        a = L.oracle_polygon_area(_ptr(xyz, F32P), n_points, None, 0, _ptr(cn, F32P))
    else:
        idx = _i32(indices)
        a = L.oracle_polygon_area(_ptr(xyz, F32P), n_points, _ptr(idx, I32P),
                                  idx.size, _ptr(cn, F32P))
    return float(a), cn


def segment_planar(xyz, normals, rows, cols, seeds, labels, cfg=None, id_offset=0,
                   robot_q=(0, 0, 0, 1)):
    L = lib()
    xyz = _f32(xyz).reshape(-1)
    normals = _f32(normals).reshape(-1)
    seeds = _i32(seeds)
    labels = _i32(labels).reshape(-1).copy()
    c = PlanarConfig(**(cfg or default_planar_config()))
    q = (C.c_float * 4)(*[float(v) for v in robot_q])
    h = C.c_void_p()
    st = L.oracle_segment_planar(_ptr(xyz, F32P), _ptr(normals, F32P), rows, cols,
                                 q, _ptr(seeds, I32P), seeds.size, C.byref(c),
                                 id_offset, _ptr(labels, I32P), C.byref(h))
    if st != 0:
        raise ValueError("oracle_segment_planar status %d" % st)
    r = collect_regions(L, "oracle_", h, "planar")
    L.oracle_regions_free(h)
    return labels, r


def segment_cluster(xyz, rows, cols, seeds, labels, cfg=None, id_offset=0):
    L = lib()
    xyz = _f32(xyz).reshape(-1)
    seeds = _i32(seeds)
    labels = _i32(labels).reshape(-1).copy()
    c = ClusterConfig(**(cfg or default_cluster_config()))
    h = C.c_void_p()
    st = L.oracle_segment_cluster(_ptr(xyz, F32P), rows, cols, _ptr(seeds, I32P),
                                  seeds.size, C.byref(c), id_offset,
                                  _ptr(labels, I32P), C.byref(h))
    if st != 0:
        raise ValueError("oracle_segment_cluster status %d" % st)
    r = collect_regions(L, "oracle_", h, "cluster")
    L.oracle_regions_free(h)
    return labels, r


def mean_shift(xyz, rows, cols, labels, cfg=None, iterations=5, id_offset=0,
               sort_mode=0):
    L = lib()
    xyz = _f32(xyz).reshape(-1)
    labels = _i32(labels).reshape(-1).copy()
    c = ClusterConfig(**(cfg or default_cluster_config()))
    h = C.c_void_p()
    modes = np.zeros(rows * cols * 8, np.float32)
    nm = C.c_int(0)
    st = L.oracle_mean_shift(_ptr(xyz, F32P), rows, cols, C.byref(c), iterations,
                             id_offset, sort_mode, _ptr(labels, I32P), C.byref(h),
                             _ptr(modes, F32P), rows * cols, C.byref(nm))
    if st != 0:
        raise ValueError("oracle_mean_shift status %d" % st)
    r = collect_regions(L, "oracle_", h, "meanshift")
    L.oracle_regions_free(h)
    return labels, r, modes.reshape(-1, 8)[:nm.value].copy()


def cluster_unorganized(xyz, labels, radius=0.5, min_region_inliers=10, id_offset=0):
    """Voxel-grid Euclidean clustering of an unorganized cloud (extension; csrc/voxel_cluster.cuh)."""
    L = lib()
    xyz = _f32(xyz).reshape(-1, 3)
    lab = _i32(labels).reshape(-1).copy()
    h = C.c_void_p()
    L.oracle_cluster_unorganized.restype = C.c_int
    L.oracle_regions_bboxes.restype = F32P
    L.oracle_regions_bboxes.argtypes = [C.c_void_p]
    st = L.oracle_cluster_unorganized(_ptr(xyz, F32P), C.c_int(xyz.shape[0]), C.c_float(radius),
                                      C.c_int(min_region_inliers), C.c_int(id_offset), _ptr(lab, I32P), C.byref(h))
    assert st == 0, st
    r = collect_regions(L, "oracle_", h, "unorganized")
    L.oracle_regions_free(h)
    return lab, r


def seeds_plane_support(xyz, normals, rows, cols, neighborhood_size=9,
                        max_plane_distance=0.05, min_support=12):
    L = lib()
    xyz = _f32(xyz).reshape(-1)
    normals = _f32(normals).reshape(-1)
    out = np.zeros(rows * cols, np.int32)
    n = L.oracle_seeds_plane_support(_ptr(xyz, F32P), _ptr(normals, F32P), rows,
                                     cols, neighborhood_size,
                                     C.c_float(max_plane_distance), min_support,
                                     _ptr(out, I32P), out.size)
    if n < 0:
        raise ValueError("oracle_seeds_plane_support status %d" % n)
    return out[:n].copy()


def seeds_average_normals(xyz, normals, rows, cols, neighborhood_size=5,
                          min_num_valid_normals=8, min_avg_normal_length=0.9999):
    L = lib()
    xyz = _f32(xyz).reshape(-1)
    normals = _f32(normals).reshape(-1)
    out = np.zeros(rows * cols, np.int32)
    n = L.oracle_seeds_average_normals(_ptr(xyz, F32P), _ptr(normals, F32P), rows,
                                       cols, neighborhood_size,
                                       min_num_valid_normals,
                                       C.c_float(min_avg_normal_length),
                                       _ptr(out, I32P), out.size)
    if n < 0:
        raise ValueError("oracle_seeds_average_normals status %d" % n)
    return out[:n].copy()


def port_lib():
    """The restatement (liboracle.so) even inside `with reference():` -- for functions the
    shim-compiled reference build does not export."""
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(so):
            build()
        _LIB = _bind(so)
    return _LIB


def transform_in_place(points, normals, q_xyzw, translation):
    """ps/multichannel_cloud.cc:53-68; returns transformed copies (either may be None)."""
    L = port_lib()
    p = _f32(points).copy() if points is not None else None
    n = _f32(normals).copy() if normals is not None else None
    cnt = (p if p is not None else n).reshape(-1, 3).shape[0]
    q = (C.c_double * 4)(*[float(v) for v in q_xyzw])
    t = (C.c_double * 3)(*[float(v) for v in translation])
    L.oracle_transform_in_place(_ptr(p, F32P) if p is not None else None,
                                _ptr(n, F32P) if n is not None else None, cnt, q, t)
    return p, n
