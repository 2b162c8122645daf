"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle
on the same seeded inputs.  Labels, inlier index sets, boundaries: bit-exact.
Normals, planes, centroids: bit-exact expected (same fp32 op order, trig
evaluated in double and rounded once on both sides); asserted to 1e-5 relative
as BASELINE.json's north_star states, and the exact-match rate is reported."""
import numpy as np
import pytest

import oracle
import pcseg
import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = pcseg.Context(0)
    yield c
    c.close()


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def assert_same_float(a, b, what, rtol=1e-5):
    a = np.asarray(a, np.float32)
    b = np.asarray(b, np.float32)
    assert a.shape == b.shape, what
    assert np.array_equal(np.isnan(a), np.isnan(b)), what + ": NaN pattern"
    m = ~np.isnan(a)
    exact = (bits(a)[m] == bits(b)[m]).mean() if m.any() else 1.0
    scale = max(float(np.abs(b[m]).max()) if m.any() else 1.0, 1e-30)
    assert np.all(np.abs(a[m] - b[m]) <= rtol * scale), what
    return exact


def assert_regions_equal(g, o, kind):
    assert g.n == o.n
    assert np.array_equal(g.inlier_offsets, o.inlier_offsets)
    assert np.array_equal(g.inlier_indices, o.inlier_indices)
    assert np.array_equal(g.label_ids, o.label_ids)
    if kind == "planar":
        assert np.array_equal(g.seed_indices, o.seed_indices)
        assert np.array_equal(g.boundary_offsets, o.boundary_offsets)
        assert np.array_equal(g.boundary_indices, o.boundary_indices)
        assert np.array_equal(g.disc_offsets, o.disc_offsets)
        # the discontinuity output is a set (absl::flat_hash_set, ps/planar_region.h:462): the oracle lists it
        # sorted by cloud index, the CUDA path in inlier order (what CreatePlanarDetectedObjectProto walks)
        for k in range(g.n):
            a, b = g.disc_offsets[k], g.disc_offsets[k + 1]
            assert np.array_equal(np.sort(g.disc_indices[a:b]), np.sort(o.disc_indices[a:b]))
        assert_same_float(g.planes, o.planes, "planes")
        assert_same_float(g.centroids, o.centroids, "centroids")
        assert_same_float(g.areas, o.areas, "areas")
        assert_same_float(g.estimators, o.estimators, "estimators")
        assert np.array_equal(g.hull_offsets, o.hull_offsets)
        assert_same_float(g.hull_points, o.hull_points, "hull")


def test_clustering_golden(ctx):
    """ps/clustering_test.cc:115-151 through the CUDA path."""
    from test_oracle_golden import CLUSTERS_CHECK
    pts, seeds = synth.clustering_test_cloud()
    labels = ctx.prepare_unlabeled_labels(100)
    out, regs = ctx.segment_cluster(pts, 10, 10, seeds, labels, id_offset=1)
    assert np.array_equal(out.reshape(10, 10).T, CLUSTERS_CHECK)
    assert [len(x) for x in regs.inliers] == [16, 20, 10, 10, 10, 27]
    o_out, o_regs = oracle.segment_cluster(pts, 10, 10, seeds, labels, id_offset=1)
    assert np.array_equal(out, o_out)
    assert_regions_equal(regs, o_regs, "cluster")


def _frame(rows, cols, seed=1234, **kw):
    return synth.depth_frame(rows, cols, seed=seed, **kw)


@pytest.mark.parametrize("rows,cols,min_d", [(60, 80, 0.1), (120, 160, 0.1), (120, 160, 0.02),
                                             (480, 640, 0.1)])
def test_normals_parity(ctx, rows, cols, min_d):
    xyz = _frame(rows, cols)
    params = dict(min_neighbor_distance=min_d)
    g, gc = ctx.normals_organized(xyz, rows, cols, params=params, want_curvature=True)
    o, oc = oracle.normals_organized(xyz, rows, cols, params=oracle.default_normals_params(**params),
                                     want_curvature=True)
    ex = assert_same_float(g, o, "normals")
    assert_same_float(gc, oc, "curvature")
    assert ex == 1.0, "normals exact-match rate %.6f" % ex


def test_normals_subrange_and_no_diag(ctx):
    rows, cols = 96, 128
    xyz = _frame(rows, cols, seed=7)
    init = np.full((rows * cols, 3), 7.0, np.float32)
    params = dict(include_diagonal_neighbors=0, min_num_support_neighbors=3, min_neighbor_distance=0.05)
    g = ctx.normals_organized(xyz, rows, cols, sensor_t=(0.1, -0.2, 0.3), params=params,
                              row_range=(5, 90), col_range=(3, 100), normals_init=init)
    o = oracle.normals_organized(xyz, rows, cols, sensor_t=(0.1, -0.2, 0.3),
                                 params=oracle.default_normals_params(**params),
                                 row_range=(5, 90), col_range=(3, 100), normals_init=init)
    assert assert_same_float(g, o, "normals subrange") == 1.0


def test_normals_invalid_plane_chain(ctx):
    """Quirk Q3: estimator state leaks after an invalid plane (collinear support)."""
    rows, cols = 16, 24
    xyz = np.zeros((rows * cols, 3), np.float32)
    rng = np.random.default_rng(5)
    for c in range(cols):
        for r in range(rows):
            xyz[c * rows + r] = (0.2 * c, 0.2 * r, 1.0 + 0.01 * rng.standard_normal())
    # a block of exactly collinear / identical points -> invalid planes
    for c in range(6, 12):
        for r in range(4, 10):
            xyz[c * rows + r] = (1.0, 1.0, 1.0)
    params = dict(min_neighbor_distance=0.0, max_neighbor_distance=10.0)
    g = ctx.normals_organized(xyz, rows, cols, params=params)
    o = oracle.normals_organized(xyz, rows, cols, params=oracle.default_normals_params(**params))
    assert assert_same_float(g, o, "normals chain") == 1.0


def _planar_case(ctx, xyz, rows, cols, seeds, cfg=None, nparams=None):
    nparams = nparams or {}
    nrm = oracle.normals_organized(xyz, rows, cols, params=oracle.default_normals_params(**nparams))
    labels = np.full(rows * cols, -1, np.int32)
    o_lab, o_regs = oracle.segment_planar(xyz, nrm, rows, cols, seeds, labels,
                                          cfg=oracle.default_planar_config(**(cfg or {})))
    g_lab, g_regs = ctx.segment_planar(xyz, nrm, rows, cols, seeds, labels, cfg=cfg)
    assert np.array_equal(g_lab, o_lab), "labels differ at %d cells" % (g_lab != o_lab).sum()
    assert_regions_equal(g_regs, o_regs, "planar")
    return g_lab, g_regs, nrm


@pytest.mark.parametrize("rows,cols", [(60, 80), (120, 160), (240, 320)])
def test_planar_parity_small(ctx, rows, cols):
    xyz = _frame(rows, cols)
    seeds = np.arange(rows * cols - 1, -1, -1, dtype=np.int32)
    _, regs, _ = _planar_case(ctx, xyz, rows, cols, seeds, nparams=dict(min_neighbor_distance=0.05))
    assert regs.n >= 1


@pytest.mark.parametrize("mode", ["global", "pipe", "wave1"])
def test_planar_alternative_paths(ctx, monkeypatch, mode):
    """The older planar growth paths stay selectable (PCSEG_PLANAR_MODE, read per call): everything through
    L2 (also the fallback for clouds whose mask does not fit shared memory), the warp-specialised
    32-entry pipeline of round 1, and the first window-parallel version."""
    monkeypatch.setenv("PCSEG_PLANAR_MODE", mode)
    rows, cols = 120, 160
    xyz = _frame(rows, cols, seed=5, yaw_deg=-8.0)
    seeds = np.arange(rows * cols - 1, -1, -1, dtype=np.int32)
    for cfg in (None, dict(plane_model_reestimation_period=7, max_plane_distance=0.02)):
        _, regs, _ = _planar_case(ctx, xyz, rows, cols, seeds, cfg=cfg, nparams=dict(min_neighbor_distance=0.05))
        assert regs.n >= 1


def test_planar_above_shared_memory_mask(ctx):
    """A cloud of 2.4 M cells: its 1-bit mask (300 KB) does not fit shared memory, so the planes are grown
    by the L2 path without any environment override.  Two tilted sheets with a step between them."""
    rows, cols = 1200, 2000
    c, r = np.meshgrid(np.arange(cols, dtype=np.float32), np.arange(rows, dtype=np.float32), indexing="ij")
    z = np.where(c < cols // 2, 2.0, 2.8).astype(np.float32) + 0.0005 * r
    xyz = np.stack([0.004 * c, 0.004 * r, z], axis=-1).reshape(-1, 3).astype(np.float32)
    nrm = np.zeros_like(xyz)
    nrm[:, 2] = -1.0
    seeds = np.array([10 * rows + 10, (cols - 10) * rows + 10], np.int32)
    labels = np.full(rows * cols, -1, np.int32)
    cfg = dict(plane_model_reestimation_period=200)
    o_lab, o_r = oracle.segment_planar(xyz, nrm, rows, cols, seeds, labels, cfg=oracle.default_planar_config(**cfg))
    g_lab, g_r = ctx.segment_planar(xyz, nrm, rows, cols, seeds, labels, cfg=cfg)
    assert o_r.n == 2
    assert np.array_equal(g_lab, o_lab)
    assert_regions_equal(g_r, o_r, "planar")


def test_planar_parity_configs(ctx):
    rows, cols = 120, 160
    xyz = _frame(rows, cols, seed=77, yaw_deg=12.0)
    seeds = np.arange(rows * cols - 1, -1, -1, dtype=np.int32)
    for cfg in (dict(plane_model_reestimation_period=7, max_plane_distance=0.02),
                dict(plane_model_reestimation_period=1, min_region_inliers=50),
                dict(plane_model_reestimation_period=64, min_region_area=0.5),
                dict(plane_model_reestimation_period=30, max_plane_distance=0.005, min_region_inliers=3)):
        _planar_case(ctx, xyz, rows, cols, seeds, cfg=cfg, nparams=dict(min_neighbor_distance=0.05))


def test_planar_seed_orders(ctx):
    rows, cols = 120, 160
    xyz = _frame(rows, cols, seed=3)
    rng = np.random.default_rng(1)
    seeds = rng.permutation(rows * cols).astype(np.int32)[:5000]
    _planar_case(ctx, xyz, rows, cols, seeds, nparams=dict(min_neighbor_distance=0.05))
    nrm = oracle.normals_organized(xyz, rows, cols)
    ps = oracle.seeds_plane_support(xyz, nrm, rows, cols)
    _planar_case(ctx, xyz, rows, cols, ps)
    _planar_case(ctx, xyz, rows, cols, np.zeros(0, np.int32))


@pytest.mark.parametrize("checker", ["port", "reference"])
def test_c1_full_pipeline_parity(ctx, checker):
    """Config C1: 640x480, normals -> planes -> clusters, all stages vs the oracle: the CPU
    restatement and the reference's own sources compiled in place (oracle/_ref)."""
    import contextlib
    if checker == "reference" and not oracle.ref_available():
        pytest.skip("oracle/_ref/libref.so not built on this machine")
    with (oracle.reference() if checker == "reference" else contextlib.nullcontext()):
        _c1_full_pipeline(ctx)


def _c1_full_pipeline(ctx):
    rows, cols = 480, 640
    xyz = synth.c1()
    g_n = ctx.normals_organized(xyz, rows, cols)
    o_n = oracle.normals_organized(xyz, rows, cols)
    assert assert_same_float(g_n, o_n, "normals") == 1.0
    seeds = np.arange(rows * cols - 1, -1, -1, dtype=np.int32)
    labels = np.full(rows * cols, -1, np.int32)
    o_lab, o_regs = oracle.segment_planar(xyz, o_n, rows, cols, seeds, labels)
    g_lab, g_regs = ctx.segment_planar(xyz, g_n, rows, cols, seeds, labels)
    assert np.array_equal(g_lab, o_lab)
    assert_regions_equal(g_regs, o_regs, "planar")
    assert g_regs.n >= 4
    o_lab2, o_c = oracle.segment_cluster(xyz, rows, cols, seeds, o_lab, id_offset=o_regs.n)
    g_lab2, g_c = ctx.segment_cluster(xyz, rows, cols, seeds, g_lab, id_offset=g_regs.n)
    assert np.array_equal(g_lab2, o_lab2)
    assert_regions_equal(g_c, o_c, "cluster")
    # second seed variant of SURVEY 8d: FindSeedPointsFromPlaneSupport seeds
    ps_o = oracle.seeds_plane_support(xyz, o_n, rows, cols)
    ps_g = ctx.seeds_plane_support(xyz, g_n, rows, cols)
    assert np.array_equal(ps_g, ps_o)
    o_lab3, o_r3 = oracle.segment_planar(xyz, o_n, rows, cols, ps_o, labels)
    g_lab3, g_r3 = ctx.segment_planar(xyz, g_n, rows, cols, ps_g, labels)
    assert np.array_equal(g_lab3, o_lab3)
    assert_regions_equal(g_r3, o_r3, "planar")


def test_cluster_parity_lidar_like(ctx):
    """C4-style: range image with masked ground, half window 2, threshold 0.25."""
    xyz, labels = synth.c4(rows=32, cols=1024, n_objects=60)
    rows, cols = 32, 1024
    seeds = np.arange(rows * cols - 1, -1, -1, dtype=np.int32)
    cfg = dict(squared_distance_threshold=0.25, half_search_window=2, min_region_inliers=10)
    o_lab, o_r = oracle.segment_cluster(xyz, rows, cols, seeds, labels, cfg=oracle.default_cluster_config(**cfg))
    g_lab, g_r = ctx.segment_cluster(xyz, rows, cols, seeds, labels, cfg=cfg)
    assert np.array_equal(g_lab, o_lab)
    assert_regions_equal(g_r, o_r, "cluster")
    assert g_r.n >= 3


def test_cluster_parity_noisy_thresholds(ctx):
    """Exercises quirk Q5 (reject / re-enqueue) with a tight threshold on noise."""
    rows, cols = 64, 96
    rng = np.random.default_rng(11)
    xyz = np.zeros((rows * cols, 3), np.float32)
    for c in range(cols):
        for r in range(rows):
            xyz[c * rows + r] = (0.05 * c, 0.05 * r, 0.08 * rng.standard_normal())
    xyz[rng.random(rows * cols) < 0.05] = np.nan
    seeds = rng.permutation(rows * cols).astype(np.int32)
    labels = np.full(rows * cols, -1, np.int32)
    labels[rng.random(rows * cols) < 0.03] = -2
    for h, thr, mi in ((1, 0.012, 7), (2, 0.02, 3), (3, 0.008, 1), (4, 0.03, 2), (5, 0.012, 4), (7, 0.02, 1)):
        cfg = dict(squared_distance_threshold=thr, half_search_window=h, min_region_inliers=mi)
        o_lab, o_r = oracle.segment_cluster(xyz, rows, cols, seeds, labels, cfg=oracle.default_cluster_config(**cfg))
        g_lab, g_r = ctx.segment_cluster(xyz, rows, cols, seeds, labels, cfg=cfg)
        assert np.array_equal(g_lab, o_lab), (h, thr, mi)
        assert_regions_equal(g_r, o_r, "cluster")


def test_cluster_shared_memory_path(ctx, monkeypatch):
    """The opt-in shared-memory cluster path (grow_cluster_wave.cuh, PCSEG_CLUSTER_MODE=wave; h == 1) on the
    reference's golden cloud, on the noisy Q5 case and on a depth frame after the planar stage."""
    monkeypatch.setenv("PCSEG_CLUSTER_MODE", "wave")
    test_clustering_golden(ctx)
    rows, cols = 64, 96
    rng = np.random.default_rng(11)
    xyz = np.zeros((rows * cols, 3), np.float32)
    xyz[:, 0] = np.repeat(0.05 * np.arange(cols), rows)
    xyz[:, 1] = np.tile(0.05 * np.arange(rows), cols)
    xyz[:, 2] = 0.08 * rng.standard_normal(rows * cols)
    xyz[rng.random(rows * cols) < 0.05] = np.nan
    seeds = rng.permutation(rows * cols).astype(np.int32)
    labels = np.full(rows * cols, -1, np.int32)
    labels[rng.random(rows * cols) < 0.03] = -2
    for thr, mi in ((0.012, 7), (0.02, 1), (0.006, 2)):
        cfg = dict(squared_distance_threshold=thr, half_search_window=1, min_region_inliers=mi)
        o_lab, o_r = oracle.segment_cluster(xyz, rows, cols, seeds, labels, cfg=oracle.default_cluster_config(**cfg))
        g_lab, g_r = ctx.segment_cluster(xyz, rows, cols, seeds, labels, cfg=cfg)
        assert np.array_equal(g_lab, o_lab), (thr, mi)
        assert_regions_equal(g_r, o_r, "cluster")
    rows, cols = 240, 320
    xyz = _frame(rows, cols, seed=77)
    nrm = oracle.normals_organized(xyz, rows, cols, params=oracle.default_normals_params(min_neighbor_distance=0.04))
    alld = np.arange(rows * cols - 1, -1, -1, dtype=np.int32)
    o_lab, o_r = oracle.segment_planar(xyz, nrm, rows, cols, alld, np.full(rows * cols, -1, np.int32))
    o_lab2, o_c = oracle.segment_cluster(xyz, rows, cols, alld, o_lab, id_offset=o_r.n)
    g_lab2, g_c = ctx.segment_cluster(xyz, rows, cols, alld, o_lab, id_offset=o_r.n)
    assert np.array_equal(g_lab2, o_lab2)
    assert_regions_equal(g_c, o_c, "cluster")


def test_mean_shift_modes_parity(ctx):
    rows, cols = 200, 300
    xyz = synth.c5(rows=rows, cols=cols, n_blobs=40)
    labels = np.full(rows * cols, -1, np.int32)
    labels[np.random.default_rng(2).random(rows * cols) < 0.02] = 0
    g = ctx.mean_shift_modes(xyz, rows, cols, labels, iterations=5)
    _, _, o = oracle.mean_shift(xyz, rows, cols, labels, iterations=5)
    seed_cells = o[:, 7].astype(np.int64)
    assert np.array_equal(np.nonzero(g[:, 7] == 1.0)[0], seed_cells)
    gs = g[seed_cells]
    assert np.array_equal(gs[:, 6], o[:, 6])
    # modes within 1e-4 of the bandwidth (1.0) as the north star states; bit-exact expected
    assert np.all(np.abs(gs[:, :5] - o[:, :5]) <= 1e-4)
    assert np.array_equal(bits(gs[:, :6]), bits(o[:, :6]))


def _assert_ms_equal(g_lab, g_r, o_lab, o_r):
    assert np.array_equal(g_lab, o_lab)
    assert_regions_equal(g_r, o_r, "cluster")
    assert np.all(np.abs(g_r.seed_points - o_r.seed_points) <= 1e-4)  # 1e-4 of the 1 m bandwidth
    assert np.array_equal(bits(g_r.seed_points), bits(o_r.seed_points))


@pytest.mark.parametrize("rows,cols,pitch,blobs,h,sort_mode,min_inl", [
    (90, 120, 0.12, 8, 1, 0, 7), (90, 120, 0.12, 8, 2, 1, 7), (200, 300, 0.05, 40, 1, 0, 7),
    (120, 90, 0.2, 10, 3, 0, 50), (64, 64, 0.3, 6, 1, 0, 400)])
def test_mean_shift_full_parity(ctx, rows, cols, pitch, blobs, h, sort_mode, min_inl):
    """SlidingMeanShift end to end (iteration + sort + growth), ps/mean_shift_segmentation.h:207-330."""
    xyz = synth.c5(rows=rows, cols=cols, n_blobs=blobs, pitch=pitch)
    labels = np.full(rows * cols, -1, np.int32)
    labels[np.random.default_rng(3).random(rows * cols) < 0.03] = -2  # masked cells are never seeds
    cfg = dict(half_search_window=h, min_region_inliers=min_inl)
    o_lab, o_r, _ = oracle.mean_shift(xyz, rows, cols, labels, cfg=oracle.default_cluster_config(**cfg),
                                      iterations=4, id_offset=3, sort_mode=sort_mode)
    g_lab, g_r = ctx.mean_shift(xyz, rows, cols, labels, cfg=cfg, iterations=4, id_offset=3, sort_mode=sort_mode)
    _assert_ms_equal(g_lab, g_r, o_lab, o_r)


@pytest.mark.parametrize("rows,cols,h,thr,min_inl", [(512, 520, 1, 0.004, 7), (128, 2100, 2, 0.01, 10), (300, 900, 3, 0.02, 25),
                                                     (270, 1000, 5, 0.03, 40), (512, 513, 1, 0.004, 1)])
def test_cluster_single_large_frame_paths(ctx, rows, cols, h, thr, min_inl):
    """pcseg_segment_cluster on one frame of >= 2^18 cells with the all-descending seed list: one CTA per
    connected component (default), column bands only (PCSEG_CLUSTER_COMPONENTS=0) and one CTA for the
    frame (PCSEG_CLUSTER_BANDS=0) all equal the oracle."""
    import os
    assert rows * cols >= 1 << 18
    xyz = synth.speckle_cloud(rows, cols, seed=rows + h)
    n = rows * cols
    labels = np.full(n, -1, np.int32)
    labels[np.random.default_rng(5).random(n) < 0.02] = 3      # cells a planar stage took
    seeds = np.arange(n - 1, -1, -1, dtype=np.int32)
    cfg = dict(squared_distance_threshold=thr, half_search_window=h, min_region_inliers=min_inl)
    o_lab, o_r = oracle.segment_cluster(xyz, rows, cols, seeds, labels, cfg=oracle.default_cluster_config(**cfg), id_offset=4)
    assert o_r.n >= 20 and int((o_lab == -1).sum()) > 100
    for env in ({}, {"PCSEG_CLUSTER_COMPONENTS": "0"}, {"PCSEG_CLUSTER_BANDS": "0"}):
        os.environ.update(env)
        try:
            g_lab, g_r = ctx.segment_cluster(xyz, rows, cols, seeds, labels, cfg=cfg, id_offset=4)
        finally:
            for k in env:
                del os.environ[k]
        assert np.array_equal(g_lab, o_lab), env
        assert_regions_equal(g_r, o_r, "cluster")


def test_cluster_single_large_frame_degenerate(ctx):
    """One frame of >= 2^18 cells with nothing to grow (every cell labelled / every point NaN), with a single
    unlabeled cell, and with one component only: the component path equals the oracle."""
    rows, cols = 512, 520
    n = rows * cols
    seeds = np.arange(n - 1, -1, -1, dtype=np.int32)
    xyz = synth.speckle_cloud(rows, cols, seed=77)
    cfg = dict(squared_distance_threshold=0.004, half_search_window=1, min_region_inliers=1)
    cases = []
    cases.append((xyz, np.full(n, 5, np.int32)))                                  # every cell labelled
    cases.append((np.full((n, 3), np.nan, np.float32), np.full(n, -1, np.int32)))  # every point NaN
    one = np.full(n, 5, np.int32); one[12345] = -1
    p1 = xyz.copy(); p1[12345] = (1.0, 2.0, 3.0)
    cases.append((p1, one))                                                       # a single unlabeled cell
    flat = np.zeros((n, 3), np.float32)
    cc, rr = np.meshgrid(np.arange(cols), np.arange(rows), indexing="ij")
    flat[:, 0] = cc.reshape(-1) * 0.01; flat[:, 1] = rr.reshape(-1) * 0.01
    cases.append((flat, np.full(n, -1, np.int32)))                                # one component = the frame
    for k, (pts, lab) in enumerate(cases):
        o_lab, o_r = oracle.segment_cluster(pts, rows, cols, seeds, lab, cfg=oracle.default_cluster_config(**cfg), id_offset=2)
        g_lab, g_r = ctx.segment_cluster(pts, rows, cols, seeds, lab, cfg=cfg, id_offset=2)
        assert np.array_equal(g_lab, o_lab), k
        assert_regions_equal(g_r, o_r, "cluster")
        assert o_r.n == (0, 0, 1, 1)[k]


def test_reserved_label_ranges(ctx):
    """The wave and component paths borrow label values for temporary codes: a caller grid that already
    holds such values is refused (mean shift) or takes the other path (cluster growth) -- never corrupted."""
    rows, cols = 512, 520
    n = rows * cols
    xyz = synth.speckle_cloud(rows, cols, seed=78)
    seeds = np.arange(n - 1, -1, -1, dtype=np.int32)
    labels = np.full(n, -1, np.int32)
    labels[1000:1010] = np.iinfo(np.int32).min + 5          # inside the component-code range
    cfg = dict(squared_distance_threshold=0.004, half_search_window=1, min_region_inliers=7)
    o_lab, o_r = oracle.segment_cluster(xyz, rows, cols, seeds, labels, cfg=oracle.default_cluster_config(**cfg))
    g_lab, g_r = ctx.segment_cluster(xyz, rows, cols, seeds, labels, cfg=cfg)
    assert np.array_equal(g_lab, o_lab)
    assert_regions_equal(g_r, o_r, "cluster")
    c5 = synth.c5(rows=60, cols=80, n_blobs=4, pitch=0.05)
    lab = np.full(60 * 80, -1, np.int32)
    lab[17] = 0x40000000
    with pytest.raises(pcseg.PcsegError) as e:
        ctx.mean_shift(c5, 60, 80, lab, iterations=2)
    assert e.value.status == 6 and "0x40000000" in str(e.value)
    import os
    os.environ["PCSEG_MS_GROWTH"] = "serial"                 # the serial walk has no such limit
    try:
        g_lab, g_r = ctx.mean_shift(c5, 60, 80, lab, iterations=2)
    finally:
        del os.environ["PCSEG_MS_GROWTH"]
    o_lab, o_r, _ = oracle.mean_shift(c5, 60, 80, lab, iterations=2)
    _assert_ms_equal(g_lab, g_r, o_lab, o_r)


def test_mean_shift_single_seed_and_tiny_clouds(ctx):
    """Fewer modes than a wave has ranks: 1 x 1 ... 12 x 12 clouds, and a cloud with one unlabeled cell."""
    for rows, cols in ((1, 1), (2, 3), (12, 12), (11, 40)):
        xyz = synth.c5(rows=rows, cols=cols, n_blobs=1, pitch=0.05, nan_frac=0.0)
        labels = np.full(rows * cols, -1, np.int32)
        cfg = dict(half_search_window=1, min_region_inliers=1)
        o_lab, o_r, _ = oracle.mean_shift(xyz, rows, cols, labels, cfg=oracle.default_cluster_config(**cfg), iterations=2)
        g_lab, g_r = ctx.mean_shift(xyz, rows, cols, labels, cfg=cfg, iterations=2)
        _assert_ms_equal(g_lab, g_r, o_lab, o_r)
    rows, cols = 40, 50
    xyz = synth.c5(rows=rows, cols=cols, n_blobs=3, pitch=0.05, nan_frac=0.0)
    labels = np.full(rows * cols, 7, np.int32); labels[777] = -1
    o_lab, o_r, _ = oracle.mean_shift(xyz, rows, cols, labels, iterations=3)
    g_lab, g_r = ctx.mean_shift(xyz, rows, cols, labels, iterations=3)
    _assert_ms_equal(g_lab, g_r, o_lab, o_r)


@pytest.mark.parametrize("rows,cols,tr,tc,pitch,step,h,min_inl", [
    (96, 120, 24, 30, 0.05, 0.6, 1, 7), (96, 120, 24, 30, 0.05, 0.6, 2, 7), (150, 150, 25, 25, 0.08, 0.3, 1, 20),
    (120, 160, 12, 16, 0.1, 0.7, 3, 5), (200, 200, 20, 20, 0.03, 0.5, 2, 300)])
def test_mean_shift_touching_regions(ctx, rows, cols, tr, tc, pitch, step, h, min_inl):
    """Many regions that touch one another: the wave path (several regions in flight, resolved in
    order) and the serial path (PCSEG_MS_GROWTH=serial) both equal the oracle."""
    import os
    xyz = synth.terraces(rows, cols, tr, tc, pitch, step, seed=rows + h)
    labels = np.full(rows * cols, -1, np.int32)
    cfg = dict(half_search_window=h, min_region_inliers=min_inl)
    o_lab, o_r, _ = oracle.mean_shift(xyz, rows, cols, labels, cfg=oracle.default_cluster_config(**cfg),
                                      iterations=3, id_offset=2, sort_mode=1)
    assert o_r.n >= 4
    g_lab, g_r = ctx.mean_shift(xyz, rows, cols, labels, cfg=cfg, iterations=3, id_offset=2, sort_mode=1)
    _assert_ms_equal(g_lab, g_r, o_lab, o_r)
    os.environ["PCSEG_MS_GROWTH"] = "serial"
    try:
        s_lab, s_r = ctx.mean_shift(xyz, rows, cols, labels, cfg=cfg, iterations=3, id_offset=2, sort_mode=1)
    finally:
        del os.environ["PCSEG_MS_GROWTH"]
    _assert_ms_equal(s_lab, s_r, o_lab, o_r)
    # ranks above 0 with room for 100 cells only: they outgrow their slot, stop and are grown again
    os.environ["PCSEG_MS_SLOT_CAP"] = "100"
    try:
        c_lab, c_r = ctx.mean_shift(xyz, rows, cols, labels, cfg=cfg, iterations=3, id_offset=2, sort_mode=1)
    finally:
        del os.environ["PCSEG_MS_SLOT_CAP"]
    _assert_ms_equal(c_lab, c_r, o_lab, o_r)


def test_mean_shift_sharded_ranges_equal_full(ctx):
    """Seed-cell shards (the multi-GPU path of C5) concatenate to the unsharded records, and
    the growth phase on gathered records equals the one-call path."""
    import parallel
    rows, cols = 150, 211
    xyz = synth.c5(rows=rows, cols=cols, n_blobs=30, pitch=0.08)
    labels = np.full(rows * cols, -1, np.int32)
    full = ctx.mean_shift_modes(xyz, rows, cols, labels, iterations=5)
    for world in (2, 3, 8):
        parts = [ctx.mean_shift_modes_range(xyz, rows, cols, labels, 5, b, e)
                 for b, e in parallel.cell_ranges(rows * cols, world)]
        assert np.array_equal(bits(np.concatenate(parts)), bits(full))
    a_lab, a_r = ctx.mean_shift(xyz, rows, cols, labels, iterations=5)
    b_lab, b_r = ctx.mean_shift_from_modes(xyz, rows, cols, full, labels)
    _assert_ms_equal(b_lab, b_r, a_lab, a_r)
    assert a_r.n >= 3


def test_mean_shift_empty_and_all_labeled(ctx):
    rows, cols = 20, 30
    xyz = synth.c5(rows=rows, cols=cols, n_blobs=2, pitch=0.1)
    labels = np.zeros(rows * cols, np.int32)  # nothing unlabeled: no seeds, no regions
    g_lab, g_r = ctx.mean_shift(xyz, rows, cols, labels)
    assert g_r.n == 0 and np.array_equal(g_lab, labels)


def test_error_statuses(ctx):
    pts, seeds = synth.clustering_test_cloud()
    labels = np.full(100, -1, np.int32)
    with pytest.raises(pcseg.PcsegError) as e:
        ctx.segment_cluster(pts[:10], 1, 10, seeds[:10] % 10, labels[:10])
    assert e.value.status == 2  # CHECK_GT(points.Rows(), 1)
    with pytest.raises(pcseg.PcsegError) as e:
        ctx.segment_cluster(pts, 10, 10, np.array([100], np.int32), labels)
    assert e.value.status == 7
    with pytest.raises(pcseg.PcsegError) as e:
        ctx.segment_planar(pts, pts, 10, 10, seeds, labels, cfg=dict(plane_model_reestimation_period=0))
    assert e.value.status == 3
