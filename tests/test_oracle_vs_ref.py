"""Pins the CPU restatement (oracle/pcseg_oracle.cc) against the REFERENCE's own
code: oracle/_ref/libref.so is the reference's .h/.cc compiled in place from
/root/reference against the shim headers in oracle/shims (patch P1 only).
Runs wherever the prebuilt _ref exists (this container; it also travels to the
GPU box); skipped otherwise.  Also replays the reference's golden tests through
the reference build itself, so the shim layer is pinned too."""
import numpy as np
import pytest

import oracle
import synth
from test_oracle_golden import CLUSTERS_CHECK

pytestmark = pytest.mark.skipif(not oracle.ref_available(), reason="oracle/_ref/libref.so not built")


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def both(fn, *a, **kw):
    r0 = fn(*a, **kw)
    with oracle.reference():
        r1 = fn(*a, **kw)
    return r0, r1


def assert_regions_bitwise(a, b, kind):
    assert a.n == b.n
    for k in ("inlier_offsets", "inlier_indices", "label_ids"):
        assert np.array_equal(getattr(a, k), getattr(b, k)), k
    if kind == "planar":
        for k in ("seed_indices", "boundary_offsets", "boundary_indices", "disc_offsets", "disc_indices",
                  "hull_offsets"):
            assert np.array_equal(getattr(a, k), getattr(b, k)), k
        for k in ("planes", "centroids", "areas", "estimators"):
            assert np.array_equal(bits(getattr(a, k)), bits(getattr(b, k))), k
        # hull vertices: third-party ConvexHull/PoseFromPlane restated on both sides, compare to 1e-5
        assert np.allclose(a.hull_points, b.hull_points, rtol=0, atol=1e-5 * max(1.0, float(np.abs(b.hull_points).max(initial=0))))


def test_ref_clustering_golden():
    """ps/clustering_test.cc:115-151 through the reference's own SegmentRegions<ClusterRegion>."""
    pts, seeds = synth.clustering_test_cloud()
    with oracle.reference():
        out, regs = oracle.segment_cluster(pts, 10, 10, seeds, np.full(100, -1, np.int32), id_offset=1)
    assert np.array_equal(out.reshape(10, 10).T, CLUSTERS_CHECK)
    assert [len(x) for x in regs.inliers] == [16, 20, 10, 10, 10, 27]


def test_ref_noisy_plane_golden():
    """ps/plane_estimator_test.cc:146-166 through the reference's PlaneEstimator + shim eigen-solver."""
    pts = [(0, 0, 1.1), (0, 0, 0.9), (1, 0, 1.2), (1, 0, 0.8), (1, 1, 1.1), (1, 1, 0.9), (0, 1, 1.1), (0, 1, 0.9)]
    ops = [(2, 0, 0, 1, 0)] + [(0,) + p + (0,) for p in pts] + [(3, 0, 0, 0, 0)]
    with oracle.reference():
        q = oracle.plane_estimator_replay(ops)
    valid, _, cx, cy, cz, nx, ny, nz, d, w = q[0]
    assert valid == 1.0 and w == 8.0
    assert np.allclose([cx, cy, cz], [0.5, 0.5, 1.0], atol=1e-5)
    assert np.allclose([nx, ny, nz, d], [0, 0, 1, -1], atol=1e-5)


def test_eig3_bitwise():
    rng = np.random.default_rng(7)
    for i in range(3000):
        a = rng.normal(size=(3, 3)).astype(np.float32) * np.float32(rng.choice([1e-3, 1.0, 30.0]))
        if i % 7 == 0:
            a[:, 2] = a[:, 1]  # rank deficient
        a = a @ a.T
        a6 = np.array([a[0, 0], a[1, 0], a[2, 0], a[1, 1], a[2, 1], a[2, 2]], np.float32)
        (v0, e0), (v1, e1) = both(oracle.eig3, a6)
        assert np.array_equal(bits(v0), bits(v1)) and np.array_equal(bits(e0), bits(e1))


@pytest.mark.parametrize("native", [0, 1])
@pytest.mark.parametrize("rows,cols,seed,min_d", [(96, 128, 1234, 0.05), (120, 160, 77, 0.1)])
def test_pipeline_bitwise(rows, cols, seed, min_d, native):
    oracle.set_native_libm(native)
    with oracle.reference():
        oracle.set_native_libm(native)
    try:
        xyz = synth.depth_frame(rows, cols, seed=seed)
        p = oracle.default_normals_params(min_neighbor_distance=min_d)
        n0, n1 = both(oracle.normals_organized, xyz, rows, cols, params=p)
        assert np.array_equal(np.isnan(n0), np.isnan(n1))
        assert np.array_equal(bits(n0)[~np.isnan(n0)], bits(n1)[~np.isnan(n1)])
        s0, s1 = both(oracle.seeds_plane_support, xyz, n0, rows, cols)
        assert np.array_equal(s0, s1) and len(s0) > 0
        lab = np.full(rows * cols, -1, np.int32)
        for seeds in (s0, np.arange(rows * cols - 1, -1, -1, dtype=np.int32)):
            (l0, r0), (l1, r1) = both(oracle.segment_planar, xyz, n0, rows, cols, seeds, lab)
            assert np.array_equal(l0, l1) and r0.n > 0
            assert_regions_bitwise(r0, r1, "planar")
        allseeds = np.arange(rows * cols - 1, -1, -1, dtype=np.int32)
        (c0, q0), (c1, q1) = both(oracle.segment_cluster, xyz, rows, cols, allseeds, l0, id_offset=r0.n)
        assert np.array_equal(c0, c1)
        assert_regions_bitwise(q0, q1, "cluster")
    finally:
        oracle.set_native_libm(0)
        with oracle.reference():
            oracle.set_native_libm(0)


def test_planar_config_variants_bitwise():
    rows, cols = 96, 128
    xyz = synth.depth_frame(rows, cols, seed=5)
    n = oracle.normals_organized(xyz, rows, cols, params=oracle.default_normals_params(min_neighbor_distance=0.05))
    seeds = np.arange(rows * cols - 1, -1, -1, dtype=np.int32)
    lab = np.full(rows * cols, -1, np.int32)
    for kw in (dict(plane_model_reestimation_period=7, max_plane_distance=0.02),
               dict(min_region_inliers=500, min_region_area=0.5),
               dict(plane_model_reestimation_period=1)):
        cfg = oracle.default_planar_config(**kw)
        (l0, r0), (l1, r1) = both(oracle.segment_planar, xyz, n, rows, cols, seeds, lab, cfg=cfg,
                                  robot_q=(0.1, -0.2, 0.3, 0.9))
        assert np.array_equal(l0, l1)
        assert_regions_bitwise(r0, r1, "planar")


def test_cluster_lidar_bitwise():
    rows, cols = 48, 1024
    xyz, labels = synth.c4(rows, cols, seed=3234, n_objects=120)
    seeds = np.arange(rows * cols - 1, -1, -1, dtype=np.int32)
    for h in (1, 2, 3):
        cfg = oracle.default_cluster_config(squared_distance_threshold=0.25, half_search_window=h, min_region_inliers=10)
        (l0, r0), (l1, r1) = both(oracle.segment_cluster, xyz, rows, cols, seeds, labels, cfg=cfg)
        assert np.array_equal(l0, l1) and r0.n > 10
        assert_regions_bitwise(r0, r1, "cluster")


def test_mean_shift_bitwise():
    rows, cols = 90, 120
    xyz = synth.c5(rows, cols, seed=4234, n_blobs=8, pitch=0.12)
    lab = np.full(rows * cols, -1, np.int32)
    (l0, r0, _), (l1, r1, _) = both(oracle.mean_shift, xyz, rows, cols, lab, iterations=5)
    assert np.array_equal(l0, l1) and r0.n >= 3
    assert_regions_bitwise(r0, r1, "cluster")
    assert np.array_equal(bits(r0.seed_points), bits(r1.seed_points))


def test_seeds_average_normals_and_support_neighbors():
    rows, cols = 60, 80
    xyz = synth.depth_frame(rows, cols, seed=9)
    n = oracle.normals_organized(xyz, rows, cols, params=oracle.default_normals_params(min_neighbor_distance=0.05))
    a0, a1 = both(oracle.seeds_average_normals, xyz, n, rows, cols)
    assert np.array_equal(a0, a1)
    for (c, r) in ((0, 0), (10, 10), (79, 59), (40, 0), (0, 30)):
        f0, f1 = both(oracle.find_normal_support_neighbors, xyz, rows, cols, c, r, 0.05, 1.0)
        assert np.array_equal(f0, f1)
