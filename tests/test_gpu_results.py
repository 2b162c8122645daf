"""Batched region results (pcseg_batch_submit_regions) and the result-export kernels
 against the CPU oracle: every frame's PlanarRegion / ClusterRegion records,
inlier / boundary / discontinuity lists and hull, the Create*DetectedObjectProto gathers
(ps/detected_objects.cc:21-48, .h:62-74) and MultichannelCloud::TransformInPlace
(ps/multichannel_cloud.cc:53-68)."""
import multiprocessing as mp
import os

import numpy as np
import pytest

import oracle
import pcseg
import synth
from test_gpu_parity import assert_same_float, bits

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = pcseg.Context(0)
    yield c
    c.close()


def _oracle_frame(args):
    xyz, rows, cols, use_ref = args
    import contextlib
    with (oracle.reference() if use_ref else contextlib.nullcontext()):
        n = rows * cols
        nrm = oracle.normals_organized(xyz, rows, cols)
        seeds = np.arange(n - 1, -1, -1, dtype=np.int32)
        lab, pr = oracle.segment_planar(xyz, nrm, rows, cols, seeds, np.full(n, -1, np.int32))
        lab2, cr = oracle.segment_cluster(xyz, rows, cols, seeds, lab, id_offset=pr.n)
    return lab2, pr, cr


def _check_frame(res, f, xyz, lab_g, oracle_out, check_xyz):
    lab_o, pr, cr = oracle_out
    assert np.array_equal(lab_g, lab_o), "frame %d labels" % f
    g = res.frame_regions(f, "planar")
    assert g.n == pr.n, "frame %d: %d planar regions, oracle %d" % (f, g.n, pr.n)
    assert np.array_equal(g.label_ids, pr.label_ids)
    assert np.array_equal(g.seed_indices, pr.seed_indices)
    for k in range(pr.n):
        assert np.array_equal(g.inliers[k], pr.inliers[k])
        bo = pr.boundary_indices[pr.boundary_offsets[k]:pr.boundary_offsets[k + 1]]
        assert np.array_equal(g.boundaries[k], bo)
        do = pr.disc_indices[pr.disc_offsets[k]:pr.disc_offsets[k + 1]]   # the reference's set, sorted
        assert np.array_equal(np.sort(g.disc[k]), do)
        # CreatePlanarDetectedObjectProto, ps/detected_objects.cc:39-46: positions in inlier order
        want_pos = np.nonzero(np.isin(pr.inliers[k], do))[0]
        assert np.array_equal(g.disc_positions[k], want_pos)
        assert np.array_equal(g.disc[k], pr.inliers[k][want_pos])
        ho = pr.hull_points[pr.hull_offsets[k]:pr.hull_offsets[k + 1]]
        assert g.hulls[k].shape == ho.shape
        assert_same_float(g.hulls[k], ho, "hull")
        if check_xyz:
            assert np.array_equal(bits(g.inlier_xyz[k]), bits(xyz[pr.inliers[k]]))
            assert np.array_equal(bits(g.boundary_xyz[k]), bits(xyz[bo]))
    assert_same_float(g.planes, pr.planes.reshape(-1, 4), "planes")
    assert_same_float(g.centroids, pr.centroids.reshape(-1, 3), "centroids")
    assert_same_float(g.areas, pr.areas, "areas")
    assert_same_float(g.estimators, pr.estimators.reshape(-1, 10), "estimators")
    c = res.frame_regions(f, "cluster")
    assert c.n == cr.n
    assert np.array_equal(c.label_ids, cr.label_ids)
    for k in range(cr.n):
        assert np.array_equal(c.inliers[k], cr.inliers[k])
        if check_xyz:
            assert np.array_equal(bits(c.inlier_xyz[k]), bits(xyz[cr.inliers[k]]))
        # extension output: centroid in the fixed summation order of results_pack.cuh, extent exact
        p = xyz[cr.inliers[k]].astype(np.float64)
        n = len(p)
        B = ((n + 7) // 8 + 31) // 32 * 32
        total = np.zeros(3)
        for w in range(8):
            blk = p[w * B:(w + 1) * B]
            wsum = np.zeros(3)
            for lane in range(32):
                part = blk[lane::32]
                wsum = wsum + (np.cumsum(part, axis=0)[-1] if len(part) else np.zeros(3))  # sequential adds
            total = total + wsum
        assert np.array_equal(c.centroids[k], (total / n).astype(np.float32))
        assert np.array_equal(c.bbox_min[k], xyz[cr.inliers[k]].min(axis=0))
        assert np.array_equal(c.bbox_max[k], xyz[cr.inliers[k]].max(axis=0))


def test_batch_regions_small_frames(ctx):
    """8 frames of 120x160 with every optional output against the oracle port and, for the first
    frames, against the reference's own sources (oracle/_ref)."""
    rows, cols, F = 120, 160, 8
    N = rows * cols
    xyz = np.stack([synth.depth_frame(rows, cols, seed=100 + i, yaw_deg=-15.0 + 4 * i) for i in range(F)])
    b = pcseg.Batch(ctx, F, rows, cols)
    res = pcseg.BatchResults(F, N, inlier_xyz=True, boundary_xyz=True, hull=True)
    labels = np.zeros((F, N), np.int32)
    params = pcseg.pipeline_params(normals=pcseg.normals_params(), run_cluster=True)
    b.submit_regions(xyz, params, res, labels_out=labels)
    b.wait()
    assert res.s.error_flags == 0
    off = res.frame_offsets
    assert off[0, F] == res.s.n_planar and off[1, F] == res.s.n_planar_inliers
    assert off[3, F] == res.s.n_cluster and off[4, F] == res.s.n_cluster_inliers
    assert res.s.n_planar >= F
    for f in range(F):
        _check_frame(res, f, xyz[f], labels[f], _oracle_frame((xyz[f], rows, cols, False)), True)
    if oracle.ref_available():
        for f in range(2):
            _check_frame(res, f, xyz[f], labels[f], _oracle_frame((xyz[f], rows, cols, True)), True)
    # the single-frame accessor path goes through the same pack kernels
    r0 = b.regions(0, "planar")
    g0 = res.frame_regions(0, "planar")
    assert r0.n == g0.n and all(np.array_equal(r0.inliers[k], g0.inliers[k]) for k in range(g0.n))
    b.close()


def test_batch_regions_c2_256_frames(ctx):
    """The benchmark's batch: all 256 frames' regions of one 640x480 C2 batch against the oracle
    (CPU side on all host cores)."""
    rows, cols, F = 480, 640, 256
    N = rows * cols
    with mp.get_context("fork").Pool(min(32, os.cpu_count() or 1)) as pool:
        xyz = np.stack(pool.starmap(synth.c2_frame, [(i, F, rows, cols) for i in range(F)], chunksize=4))
        want = pool.map(_oracle_frame, [(xyz[i], rows, cols, False) for i in range(F)], chunksize=2)
    b = pcseg.Batch(ctx, F, rows, cols)
    res = pcseg.BatchResults(F, N, hull=True)
    labels = np.zeros((F, N), np.int32)
    b.submit_regions(xyz, pcseg.pipeline_params(run_cluster=True), res, labels_out=labels)
    b.wait()
    assert res.s.error_flags == 0
    for f in range(F):
        _check_frame(res, f, xyz[f], labels[f], want[f], False)
    if oracle.ref_available():   # every frame again, against the reference's own sources (oracle/_ref)
        oracle.lib()
        with mp.get_context("fork").Pool(min(32, os.cpu_count() or 1)) as pool:
            want_ref = pool.map(_oracle_frame, [(xyz[i], rows, cols, True) for i in range(F)], chunksize=2)
        for f in range(F):
            _check_frame(res, f, xyz[f], labels[f], want_ref[f], False)
    b.close()


def test_batch_results_capacity_status(ctx):
    rows, cols, F = 60, 80, 3
    xyz = np.stack([synth.depth_frame(rows, cols, seed=5 + i) for i in range(F)])
    b = pcseg.Batch(ctx, F, rows, cols)
    res = pcseg.BatchResults(F, rows * cols, hull=True)
    res.s.planar_inlier_cap = 10      # too small on purpose
    params = pcseg.pipeline_params(normals=pcseg.normals_params(min_neighbor_distance=0.05))
    b.submit_regions(xyz, params, res)
    with pytest.raises(pcseg.PcsegError) as e:
        b.wait()
    assert e.value.status == 5
    assert res.s.n_planar_inliers > 10 and res.s.n_planar >= 1   # totals say how much is needed
    b.close()


def _tile_cloud(rows, cols, tile):
    """Flat tiles of tile x tile cells at alternating heights: one plane per tile."""
    xyz = np.zeros((rows * cols, 3), np.float32)
    c, r = np.meshgrid(np.arange(cols), np.arange(rows), indexing="ij")
    z = 2.0 + 0.3 * (((c // tile) + (r // tile)) % 2) + 0.001 * ((c // tile) * 7 + (r // tile) * 13 % 11)
    xyz[:, 0] = (0.02 * c).reshape(-1)
    xyz[:, 1] = (0.02 * r).reshape(-1)
    xyz[:, 2] = z.reshape(-1)
    return xyz


def test_more_planes_than_the_initial_region_table(ctx):
    """> 4096 accepted planes in one frame (the reference has no region limit): the region table
    grows and the growth is repeated, single-frame and batched."""
    rows, cols, tile = 480, 640, 6
    N = rows * cols
    xyz = _tile_cloud(rows, cols, tile)
    nrm = np.tile(np.array([0, 0, 1], np.float32), (N, 1))
    seeds = np.arange(N - 1, -1, -1, dtype=np.int32)
    cfg = dict(min_region_area=1e-6, max_plane_distance=0.01)
    labels = np.full(N, -1, np.int32)
    o_lab, o_r = oracle.segment_planar(xyz, nrm, rows, cols, seeds, labels, cfg=oracle.default_planar_config(**cfg))
    assert o_r.n > 4096
    g_lab, g_r = ctx.segment_planar(xyz, nrm, rows, cols, seeds, labels, cfg=cfg)
    assert g_r.n == o_r.n
    assert np.array_equal(g_lab, o_lab)
    assert np.array_equal(g_r.inlier_indices, o_r.inlier_indices)
    assert np.array_equal(g_r.label_ids, o_r.label_ids)


def test_transform_in_place(ctx):
    rng = np.random.default_rng(9)
    n = 10007
    pts = (rng.standard_normal((n, 3)) * 3).astype(np.float32)
    nrm = rng.standard_normal((n, 3)).astype(np.float32)
    pts[5] = np.nan
    q = rng.standard_normal(4)
    q /= np.linalg.norm(q)
    t = rng.standard_normal(3)
    want_p, want_n = oracle.transform_in_place(pts, nrm, q, t)
    gp, gn = pts.copy(), nrm.copy()
    ctx.transform_in_place(gp, gn, q, t)
    assert assert_same_float(gp, want_p, "points") == 1.0     # bit-exact outside the NaN point
    assert assert_same_float(gn, want_n, "normals") == 1.0
    gp2 = pts.copy()
    ctx.transform_in_place(points=gp2, q_xyzw=q, translation=t)
    assert assert_same_float(gp2, want_p, "points only") == 1.0


def test_gather_points(ctx):
    rng = np.random.default_rng(10)
    xyz = rng.standard_normal((5000, 3)).astype(np.float32)
    idx = rng.permutation(5000)[:1234].astype(np.int32)
    disc = rng.choice(idx, 77, replace=False).astype(np.int32)
    out, pos = ctx.gather_points(xyz, idx, disc)
    assert np.array_equal(bits(out), bits(xyz[idx]))
    assert np.array_equal(pos, np.nonzero(np.isin(idx, disc))[0])   # ps/detected_objects.cc:41-45
    with pytest.raises(pcseg.PcsegError):
        ctx.gather_points(xyz, np.array([5000], np.int32))
