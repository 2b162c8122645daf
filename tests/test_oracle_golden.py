"""Pins the CPU oracle against the reference's own golden vectors
(SURVEY.md Appendix C): ps/clustering_test.cc, ps/plane_estimator_test.cc,
ps/segmentation_test.cc, ps/algorithms_test.cc."""
import math

import numpy as np
import pytest

import oracle
import synth

CLUSTERS_CHECK = np.array([  # ps/clustering_test.cc:127-137 (printed row by row)
    [1, 1, 1, 1, 3, 4, 4, 4, 4, 5],
    [1, 1, 1, 1, 3, 4, 4, 4, 5, 5],
    [1, 1, 1, 1, 3, 4, 4, 5, 5, 6],
    [1, 1, 1, 1, 3, 4, 5, 5, 6, 6],
    [-1, -1, -1, -1, 3, 5, 5, 6, 6, 6],
    [2, 2, 2, 2, 3, 5, 6, 6, 6, 6],
    [2, 2, 2, 2, 3, 6, 6, 6, -1, 6],
    [2, 2, 2, 2, 3, 6, 6, 6, -1, 6],
    [2, 2, 2, 2, 3, 6, 6, 6, -1, 6],
    [2, 2, 2, 2, 3, 6, 6, 6, 6, 6]], dtype=np.int32)


@pytest.mark.parametrize("native", [0, 1])
def test_clustering_golden(native):
    oracle.set_native_libm(native)
    pts, seeds = synth.clustering_test_cloud()
    labels = np.full(100, -1, np.int32)
    out, regs = oracle.segment_cluster(pts, 10, 10, seeds, labels, id_offset=1)
    got = out.reshape(10, 10).T  # [row, col] from col-major
    assert np.array_equal(got, CLUSTERS_CHECK)
    assert [len(x) for x in regs.inliers] == [16, 20, 10, 10, 10, 27]
    assert list(regs.label_ids) == [1, 2, 3, 4, 5, 6]
    oracle.set_native_libm(0)


ADD, ADDW, ORIENT, QUERY, CLEAR = 0, 1, 2, 3, 4
NOISY8 = [(0, 0, 1.1), (0, 0, 0.9), (1, 0, 1.2), (1, 0, 0.8), (1, 1, 1.1),
          (1, 1, 0.9), (0, 1, 1.1), (0, 1, 0.9)]


def test_plane_estimator_corner_cases():
    """ps/plane_estimator_test.cc:76-144"""
    eps = 1e-5
    ops = [(ORIENT, 0.1, 0.1, 0.98, 0), (QUERY, 0, 0, 0, 0),
           (ADD, 1, 2, 3, 0), (QUERY, 0, 0, 0, 0),
           (ADD, 1, 2, 3, 0), (ADD, 1, 2, 3, 0), (QUERY, 0, 0, 0, 0),
           (ADD, 0, 2, 3, 0), (QUERY, 0, 0, 0, 0),
           (ADD, 0, 0, 3, 0), (QUERY, 0, 0, 0, 0),
           (ORIENT, 0.1, 0.1, -0.98, 0), (QUERY, 0, 0, 0, 0)]
    q = oracle.plane_estimator_replay(ops)
    assert q[0, 0] == 0
    assert q[1, 0] == 0 and q[1, 1] == 0
    assert np.allclose(q[1, 2:5], [1, 2, 3], atol=eps)
    assert not np.isnan(q[1, 5:9]).any()
    assert q[2, 0] == 0 and q[2, 1] == 0
    assert np.allclose(q[2, 2:5], [1, 2, 3], atol=eps)
    assert np.allclose(q[2, 5:8], [0.1, 0.1, 0.98], atol=eps)
    assert q[3, 0] == 0
    assert np.allclose(q[3, 2:5], [0.75, 2, 3], atol=eps)
    assert not np.isnan(q[3, 5:9]).any()
    assert q[4, 0] == 1
    assert np.allclose(q[4, 2:5], [0.6, 1.6, 3.0], atol=eps)
    assert np.allclose(q[4, 5:9], [0, 0, 1, -3], atol=eps)
    assert np.allclose(q[5, 5:9], [0, 0, -1, 3], atol=eps)


@pytest.mark.parametrize("weighted", [False, True])
def test_plane_estimator_noisy(weighted):
    """ps/plane_estimator_test.cc:146-191"""
    ops = [(ORIENT, 0.1, 0.1, 0.98, 0)] + [(ADDW, x, y, z, 1.0) for x, y, z in NOISY8]
    tol = 1e-5
    if weighted:
        ops += [(ADDW, 4, 6, 7, 0.001), (ADDW, -6, 5, 8, 0.001)]
        tol = 0.04
    ops.append((QUERY, 0, 0, 0, 0))
    q = oracle.plane_estimator_replay(ops)[0]
    assert q[0] == 1
    assert np.allclose(q[2:5], [0.5, 0.5, 1.0], atol=tol)
    assert np.allclose(q[5:9], [0, 0, 1, -1], atol=tol)


def test_plane_estimator_plane_normal_fwd_rev():
    """ps/plane_estimator_test.cc:40-74"""
    n = np.array([3.001, 2.0001, 1.02], np.float32)
    n = n / np.float32(np.sqrt(np.float32((n * n).sum())))
    pts = []
    for ii in range(-100, 100):
        for jj in range(-100, 100):
            x = np.float32(ii / 10.0)
            y = np.float32(jj / 10.0)
            z = (np.float32(10) - n[0] * x - n[1] * y) / n[2]
            pts.append((float(x), float(y), float(z)))
    fwd = oracle.plane_estimator_replay([(ADD, *p, 0) for p in pts] + [(QUERY, 0, 0, 0, 0)])[0]
    rev = oracle.plane_estimator_replay([(ADD, *p, 0) for p in pts[::-1]] + [(QUERY, 0, 0, 0, 0)])[0]
    assert fwd[0] == 1 and rev[0] == 1
    s = 1.0 if np.dot(fwd[5:8], n) > 0 else -1.0
    assert np.allclose(s * fwd[5:8], n, atol=1e-4)
    s2 = 1.0 if np.dot(fwd[5:8], rev[5:8]) > 0 else -1.0
    assert np.allclose(fwd[5:8], s2 * rev[5:8], atol=1e-4)


def test_find_seed_points_plane_support():
    """ps/segmentation_test.cc:28-53"""
    k = 5
    pts = np.zeros((k * k, 3), np.float32)
    nrm = np.zeros((k * k, 3), np.float32)
    for col in range(k):
        for row in range(k):
            pts[col * k + row] = (np.float32(0.1) * col, np.float32(0.1) * row, 0)
            nrm[col * k + row] = (0, 0, 1)
    assert len(oracle.seeds_plane_support(pts, nrm, k, k, 3, 0.05, 6)) == k * k - 4
    assert len(oracle.seeds_plane_support(pts, nrm, k, k, 5, 0.05, 16)) == (k - 2) * (k - 2)


def test_polygon_normal_and_area():
    """ps/algorithms_test.cc:117-140"""
    pts = np.array([[0, 0, 1000], [1, 1, 1000], [1, 0, 1000]], np.float32)
    a, cn = oracle.polygon_area(pts)
    nn = cn / np.linalg.norm(cn)
    assert np.allclose(nn, [0, 0, -1], atol=1e-4) and abs(a - 0.5) < 1e-4
    a2, _ = oracle.polygon_area(pts, [0, 1, 2, 1, 2])
    assert abs(a2 - 0.5) < 1e-4
    pts[1, 1] = 0.0
    pts[1, 2] = 999.0
    a3, cn3 = oracle.polygon_area(pts)
    assert np.allclose(cn3 / np.linalg.norm(cn3), [0, -1, 0], atol=1e-4)
    assert abs(a3 - 0.5) < 1e-4


def test_find_normal_support_neighbors():
    """ps/algorithms_test.cc:217-243: BuildDensePointCloud(100,100,10,10)."""
    sx = sy = 100
    pts = np.zeros((sx * sy, 3), np.float32)
    # CloudBuffer(size_x, size_y) -> rows = size_x, cols = size_y; AtUnsafe(j, i)
    for i in range(sx):
        for j in range(sy):
            pts[i * sx + j] = (np.float32(i * 10.0 / sx), np.float32(j * 10.0 / sy), 1.0)
    idx = oracle.find_normal_support_neighbors(pts, sx, sy, 20, 20, 0.05, 0.4, True)
    assert len(idx) == 9
    for want in (21 * 100 + 20, 20 * 100 + 21, 19 * 100 + 20, 20 * 100 + 19, 20 * 100 + 20):
        assert want in idx


def test_eig3_known():
    ev, vec = oracle.eig3([2, 0, 0, 3, 0, 5])  # diag(2,3,5)
    assert np.allclose(ev, [2, 3, 5], atol=1e-5)
    assert abs(abs(vec[0, 0]) - 1) < 1e-5 and abs(abs(vec[2, 2]) - 1) < 1e-5
    rng = np.random.default_rng(0)
    for _ in range(200):
        a = rng.standard_normal((3, 3))
        s = (a @ a.T).astype(np.float32)
        ev, vec = oracle.eig3([s[0, 0], s[1, 0], s[2, 0], s[1, 1], s[2, 1], s[2, 2]])
        ref = np.linalg.eigvalsh(s.astype(np.float64))
        assert np.allclose(ev, ref, rtol=2e-4, atol=2e-4 * abs(ref).max())
        r = s.astype(np.float64) @ vec[:, 0] - ev[0] * vec[:, 0]
        assert np.linalg.norm(r) < 5e-3 * abs(ref).max()
