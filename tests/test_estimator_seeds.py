"""PlaneEstimator entry points and the two remaining seed finders (8f-1).

CPU (-m "not gpu"): the restatement against the reference's golden vectors
(ps/plane_estimator_test.cc:193-254 AdvancedPlaneMerging) and against the reference's own code
(oracle/_ref) on random inputs.  GPU (-m gpu): the CUDA entry points through the C ABI against the
oracle -- the reference build when it is present -- on the same inputs: query rows / accumulator
states / seed lists bit-exact."""
import math

import numpy as np
import pytest

import oracle
import synth

ADD, ADDW, ORIENT, QUERY, CLEAR, SELECT, MERGE, TRANSFORM, MERGE_POSE = range(9)
NOISY8 = [(0, 0, 1.1), (0, 0, 0.9), (1, 0, 1.2), (1, 0, 0.8), (1, 1, 1.1), (1, 1, 0.9), (0, 1, 1.1), (0, 1, 0.9)]


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def qrot(q, v):
    qv = np.asarray(q[:3], np.float64)
    uv = np.cross(qv, v)
    uv = uv + uv
    return v + q[3] * uv + np.cross(qv, uv)


def merging_ops():
    """ps/plane_estimator_test.cc:193-254: estimator 1 in frame 1, estimator 2 over the same points
    moved to frame 2, then estimator1.Merge(estimator2, frame1_pose_frame2)."""
    s, c = math.sin(math.pi / 4), math.cos(math.pi / 4)
    pose = (s, 0.0, 0.0, c, 1.0, 1.0, 1.0)        # AngleAxis(pi/2, x), translation (1, 1, 1)
    qi = (-s, 0.0, 0.0, c)
    ti = -qrot(qi, np.array([1.0, 1.0, 1.0]))
    inv = (qi[0], qi[1], qi[2], qi[3], ti[0], ti[1], ti[2])
    moved, _ = oracle.transform_in_place(np.array(NOISY8, np.float32), None, pose[:4], pose[4:])
    ops = [(ORIENT, .1, .1, .98, 0)] + [(ADD, *p, 0) for p in NOISY8] + [(QUERY, 0, 0, 0, 0)]
    ops += [(SELECT, 1, 0, 0, 0), (ORIENT, .1, -.98, .1, 0)] + [(ADD, *map(float, p), 0) for p in moved] + [(QUERY, 0, 0, 0, 0)]
    ops += [(SELECT, 0, 0, 0, 0), (MERGE_POSE, 1, 1, 0, 0), (QUERY, 0, 0, 0, 0)]
    return ops, [pose, inv]


def check_merging(q):
    eps = 1e-5
    for row, w in ((q[0], 8.0), (q[2], 16.0)):
        assert row[0] == 1.0 and row[9] == w
        assert np.allclose(row[2:5], [0.5, 0.5, 1.0], atol=eps)
        assert np.allclose(row[5:9], [0, 0, 1, -1], atol=eps)
    # frame 2: normal2 = R * (0, 0, 1) = (0, -1, 0); centroid2 = pose * (0.5, 0.5, 1) = (1.5, 0, 1.5)
    assert q[1][0] == 1.0
    assert np.allclose(q[1][2:5], [1.5, 0.0, 1.5], atol=eps)
    assert np.allclose(q[1][5:8], [0, -1, 0], atol=eps)


def random_ops(seed, n=300):
    rng = np.random.default_rng(seed)
    n_poses = 3
    poses = []
    for _ in range(n_poses):
        q = rng.standard_normal(4)
        q /= np.linalg.norm(q)
        poses.append(tuple(q) + tuple(rng.standard_normal(3)))
    ops = []
    weight = [0.0] * 4
    cur = 0
    for _ in range(n):
        r = rng.random()
        p = tuple(float(v) for v in (rng.standard_normal(3) * [2.0, 2.0, 0.05] + [1.0, -1.0, 3.0]).astype(np.float32))
        if r < 0.55:
            ops.append((ADD, *p, 0)); weight[cur] += 1
        elif r < 0.65:
            w = float(np.float32(rng.random() + 0.01))
            ops.append((ADDW, *p, w)); weight[cur] += w
        elif r < 0.70 and weight[cur] > 0:
            ops.append((ORIENT, *p, 0))
        elif r < 0.80 and weight[cur] > 0:
            ops.append((QUERY, 0, 0, 0, 0))
        elif r < 0.82:
            ops.append((CLEAR, 0, 0, 0, 0)); weight[cur] = 0.0
        elif r < 0.88:
            cur = int(rng.integers(4)); ops.append((SELECT, cur, 0, 0, 0))
        elif r < 0.92:
            k = (cur + 1 + int(rng.integers(3))) % 4
            ops.append((MERGE, k, 0, 0, 0)); weight[cur] += weight[k]
        elif r < 0.96 and weight[cur] > 0:
            ops.append((TRANSFORM, int(rng.integers(n_poses)), 0, 0, 0))
        else:
            k = (cur + 1 + int(rng.integers(3))) % 4
            if weight[k] > 0:
                ops.append((MERGE_POSE, k, int(rng.integers(n_poses)), 0, 0)); weight[cur] += weight[k]
    for k in range(4):
        if weight[k] > 0:
            ops += [(SELECT, k, 0, 0, 0), (QUERY, 0, 0, 0, 0)]
    return ops, poses


def csr_case(seed, n_points=5000, n_regions=40):
    rng = np.random.default_rng(seed)
    xyz = (rng.standard_normal((n_points, 3)) * [3.0, 3.0, 0.02] + [0, 0, 2.0]).astype(np.float32)
    sizes = rng.integers(0, 400, n_regions)
    sizes[3] = 0          # an empty region: the state right after Clear()
    sizes[5] = 1          # one point: invalid plane, old normal kept (quirk Q14)
    sizes[7] = 3
    offsets = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    indices = rng.integers(0, n_points, int(offsets[-1])).astype(np.int32)
    hints = rng.standard_normal((n_regions, 3)).astype(np.float32)
    weights = (rng.random(int(offsets[-1])) + 0.05).astype(np.float32)
    return xyz, offsets, indices, hints, weights


def last_regions_case(seed):
    rows, cols = 96, 128
    xyz = synth.depth_frame(rows, cols, seed=seed)
    nrm = oracle.normals_organized(xyz, rows, cols, params=oracle.default_normals_params(min_neighbor_distance=0.05))
    seeds = oracle.seeds_plane_support(xyz, nrm, rows, cols)
    _, regs = oracle.segment_planar(xyz, nrm, rows, cols, seeds, np.full(rows * cols, -1, np.int32))
    counts = np.diff(regs.inlier_offsets).astype(np.int32)
    # two regions with equal inlier counts pin the multimap's insertion order
    counts = np.concatenate([counts, counts[:1]])
    cen = np.concatenate([regs.centroids, regs.centroids[:1] + np.float32(0.01)])
    nr = np.concatenate([regs.planes[:, :3], regs.planes[:1, :3]])
    q = np.array([0.0, 0.0, math.sin(0.01), math.cos(0.01)], np.float32)
    t = np.array([0.01, -0.02, 0.005], np.float32)
    return xyz, nrm, counts, cen, nr, q, t


# ---------------------------------------------------------------- CPU
def test_oracle_advanced_plane_merging_golden():
    ops, poses = merging_ops()
    check_merging(oracle.plane_estimator_replay2(ops, poses))


@pytest.mark.skipif(not oracle.ref_available(), reason="oracle/_ref/libref.so not built")
def test_ref_advanced_plane_merging_golden_and_port_bitwise():
    ops, poses = merging_ops()
    with oracle.reference():
        q_ref = oracle.plane_estimator_replay2(ops, poses)
    check_merging(q_ref)
    assert np.array_equal(bits(q_ref), bits(oracle.plane_estimator_replay2(ops, poses)))


@pytest.mark.skipif(not oracle.ref_available(), reason="oracle/_ref/libref.so not built")
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_port_vs_ref_estimator_random_ops(seed):
    ops, poses = random_ops(seed)
    q0 = oracle.plane_estimator_replay2(ops, poses)
    with oracle.reference():
        q1 = oracle.plane_estimator_replay2(ops, poses)
    assert q0.shape[0] > 5
    assert np.array_equal(bits(q0), bits(q1))


@pytest.mark.skipif(not oracle.ref_available(), reason="oracle/_ref/libref.so not built")
def test_port_vs_ref_seeds_last_planar_regions():
    xyz, nrm, counts, cen, nr, q, t = last_regions_case(1234)
    a = oracle.seeds_last_planar_regions(xyz, nrm, counts, cen, nr, q, t, 0.3, 0.3)
    with oracle.reference():
        b = oracle.seeds_last_planar_regions(xyz, nrm, counts, cen, nr, q, t, 0.3, 0.3)
    assert len(a) >= 2 and np.array_equal(a, b)


# ---------------------------------------------------------------- CUDA path (C ABI)
@pytest.fixture(scope="module")
def ctx():
    import pcseg
    c = pcseg.Context(0)
    yield c
    c.close()


def checker():
    """The reference's own code when its build travelled with the repo, else the restatement."""
    return oracle.reference() if oracle.ref_available() else __import__("contextlib").nullcontext()


@pytest.mark.gpu
def test_cuda_plane_estimator_golden_vectors(ctx):
    """ps/plane_estimator_test.cc:76-254 driven through the CUDA Estimator."""
    from test_oracle_golden import NOISY8 as N8
    ops, poses = merging_ops()
    q = ctx.plane_estimator_replay(ops, poses)
    check_merging(q)
    with checker():
        assert np.array_equal(bits(q), bits(oracle.plane_estimator_replay2(ops, poses)))
    # NoisyPlaneEstimation / WeightedPlaneEstimation (:146-191)
    ops = [(ORIENT, .1, .1, .98, 0)] + [(ADDW, *p, 1.0) for p in N8] + [(ADDW, 4, 6, 7, .001), (ADDW, -6, 5, 8, .001), (QUERY, 0, 0, 0, 0)]
    q = ctx.plane_estimator_replay(ops)[0]
    assert q[0] == 1.0 and np.allclose(q[2:5], [.5, .5, 1.0], atol=.04) and np.allclose(q[5:9], [0, 0, 1, -1], atol=.04)
    with checker():
        assert np.array_equal(bits(q), bits(oracle.plane_estimator_replay2(ops)[0]))
    # PlaneNormal (:40-74): 40 000 points forwards and backwards, 1e-4
    rng = np.random.default_rng(7)
    pts = np.stack([rng.uniform(-50, 50, 40000), rng.uniform(-50, 50, 40000), np.full(40000, 1.0)], 1).astype(np.float32)
    off = np.array([0, 40000, 80000], np.int32)
    idx = np.concatenate([np.arange(40000), np.arange(39999, -1, -1)]).astype(np.int32)
    q = ctx.fit_planes(pts, off, idx, hints=[[0, 0, 1], [0, 0, 1]])
    for row in q:
        assert row[0] == 1.0 and np.allclose(row[5:9], [0, 0, 1, -1], atol=1e-4)
    with checker():
        assert np.array_equal(bits(q), bits(oracle.fit_planes_csr(pts, off, idx, hints=[[0, 0, 1], [0, 0, 1]])))


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_cuda_plane_estimator_random_ops(ctx, seed):
    ops, poses = random_ops(seed)
    q = ctx.plane_estimator_replay(ops, poses)
    with checker():
        want = oracle.plane_estimator_replay2(ops, poses)
    assert np.array_equal(bits(q), bits(want))


@pytest.mark.gpu
@pytest.mark.parametrize("weighted", [False, True])
def test_cuda_fit_planes_csr(ctx, weighted):
    xyz, offsets, indices, hints, weights = csr_case(11)
    w = weights if weighted else None
    q, st = ctx.fit_planes(xyz, offsets, indices, weights=w, hints=hints, want_state=True)
    with checker():
        want = oracle.fit_planes_csr(xyz, offsets, indices, hints=hints, weights=w)
    nonempty = np.diff(offsets) > 0
    assert np.array_equal(bits(q[nonempty]), bits(want[nonempty]))
    # an empty region is the state right after Clear() + SetNormalOrientation: invalid, plane = (hint, 0);
    # the reference's Centroid()/Curvature() return uninitialised members there, so only these are compared
    assert np.all(q[~nonempty][:, 0] == 0) and np.array_equal(bits(q[~nonempty][:, 5:8]), bits(hints[~nonempty]))
    assert np.all(q[~nonempty][:, 8:] == 0)
    assert np.array_equal(bits(st[:, 9]), bits(want[:, 9]))
    # Transform / Merge of the exported states against the class replayed op by op
    rng = np.random.default_rng(5)
    qq = rng.standard_normal(4); qq /= np.linalg.norm(qq)
    pose = tuple(qq) + (0.3, -0.2, 1.5)
    keep = [r for r in range(len(offsets) - 1) if offsets[r + 1] - offsets[r] >= 3][:6]
    moved = ctx.plane_estimators_transform(st[keep], pose)
    merged = ctx.plane_estimators_merge(st[keep], st[keep][::-1].copy(), poses=pose)
    for k, r in enumerate(keep):
        def adds(rr):
            sl = slice(int(offsets[rr]), int(offsets[rr + 1]))
            if weighted:
                return [(ADDW, *map(float, xyz[i]), float(wt)) for i, wt in zip(indices[sl], weights[sl])]
            return [(ADD, *map(float, xyz[i]), 0) for i in indices[sl]]
        other = keep[len(keep) - 1 - k]
        ops = adds(r) + [(TRANSFORM, 0, 0, 0, 0), (QUERY, 0, 0, 0, 0)]
        ops += [(SELECT, 1, 0, 0, 0)] + adds(other) + [(SELECT, 2, 0, 0, 0)] + adds(r) + [(MERGE_POSE, 1, 0, 0, 0), (QUERY, 0, 0, 0, 0)]
        with checker():
            want = oracle.plane_estimator_replay2(ops, [pose])
        for got_state, row in ((moved[k], want[0]), (merged[k], want[1])):
            # centroid = sum / weight and the weight itself identify the state; compare those bitwise
            assert np.array_equal(bits(got_state[9]), bits(row[9]))
            cen = (got_state[6:9] / got_state[9]).astype(np.float32)
            assert np.array_equal(bits(cen), bits(row[2:5]))


@pytest.mark.gpu
@pytest.mark.parametrize("shape,nbhd", [((96, 128), 5), ((120, 160), 7), ((64, 64), 3)])
def test_cuda_seeds_average_normals(ctx, shape, nbhd):
    rows, cols = shape
    xyz = synth.depth_frame(rows, cols, seed=77)
    nrm = oracle.normals_organized(xyz, rows, cols, params=oracle.default_normals_params(min_neighbor_distance=0.05))
    got = ctx.seeds_average_normals(xyz, nrm, rows, cols, nbhd, 8, 0.999)
    with checker():
        want = oracle.seeds_average_normals(xyz, nrm, rows, cols, nbhd, 8, 0.999)
    assert len(want) > 50
    assert np.array_equal(got, want)


@pytest.mark.gpu
def test_cuda_seeds_average_normals_5x5_plane(ctx):
    """The 5x5 plane of ps/segmentation_test.cc:28-44 (the reference only runs the plane-support finder
    on it): window 3 -> rows 1..3 x cols 2..4 = 9 seeds, index = LinearizeIndex(row, col) - 1 (Q1)."""
    rows = cols = 5
    pts = np.zeros((25, 3), np.float32)
    nrm = np.zeros((25, 3), np.float32)
    for c in range(cols):
        for r in range(rows):
            pts[c * rows + r] = (0.1 * c, 0.1 * r, 1.0)
            nrm[c * rows + r] = (0.0, 0.0, 1.0)
    got = ctx.seeds_average_normals(pts, nrm, rows, cols, 3, 8, 0.9999)
    want = oracle.seeds_average_normals(pts, nrm, rows, cols, 3, 8, 0.9999)
    assert np.array_equal(got, want) and len(got) == 9


@pytest.mark.gpu
def test_cuda_seeds_last_planar_regions(ctx):
    xyz, nrm, counts, cen, nr, q, t = last_regions_case(1234)
    got = ctx.seeds_last_planar_regions(xyz, nrm, counts, cen, nr, q, t, 0.3, 0.3, seeds=[7, 9])
    with checker():
        want = oracle.seeds_last_planar_regions(xyz, nrm, counts, cen, nr, q, t, 0.3, 0.3)
    assert len(want) >= 2
    assert np.array_equal(got, np.concatenate([[7, 9], want]))
    # nothing within reach: no seed appended
    none = ctx.seeds_last_planar_regions(xyz, nrm, counts, cen + np.float32(50.0), nr, q, t, 0.3, 0.3)
    assert none.size == 0
