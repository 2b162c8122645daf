"""Golden fixtures under tests/golden/ were produced by the REFERENCE's own code
(oracle/_ref, tools/make_golden.py).  They do not need /root/reference at test
time.  -m "not gpu": the CPU restatement must reproduce them; -m gpu: the CUDA
path must reproduce them through the C ABI.
Bars: labels / index lists bit-exact; planes, centroids, areas 1e-5 relative;
normals 1e-5 relative (bit-exact in practice); mean-shift modes 1e-4 of the bandwidth."""
import os

import numpy as np
import pytest

import oracle

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
DEPTH = ["depth_96x128_alldesc", "depth_96x128_planesupport"]


def load(name):
    return np.load(os.path.join(G, name + ".npz"))


def close(a, b, rtol=1e-5):
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    if a.shape != b.shape or not np.array_equal(np.isnan(a), np.isnan(b)):
        return False
    m = ~np.isnan(a)
    scale = max(float(np.abs(b[m]).max()) if m.any() else 1.0, 1e-30)
    return bool(np.all(np.abs(a[m] - b[m]) <= rtol * scale))


def check_regions(r, f, prefix, kind):
    assert r.n == len(f[prefix + "label_ids"])
    for k in ("inlier_offsets", "inlier_indices", "label_ids"):
        assert np.array_equal(getattr(r, k), f[prefix + k]), k
    if kind == "planar":
        for k in ("seed_indices", "boundary_offsets", "boundary_indices", "disc_offsets", "hull_offsets"):
            assert np.array_equal(getattr(r, k), f[prefix + k]), k
        # the discontinuity output is a set (absl::flat_hash_set, ps/planar_region.h:462): the fixture lists it
        # by cloud index, the CUDA path in inlier order (what CreatePlanarDetectedObjectProto walks)
        for i in range(r.n):
            a, b = r.disc_offsets[i], r.disc_offsets[i + 1]
            assert np.array_equal(np.sort(r.disc_indices[a:b]), np.sort(f[prefix + "disc_indices"][a:b])), "disc_indices"
        for k in ("planes", "centroids", "areas", "estimators", "hull_points"):
            assert close(getattr(r, k), f[prefix + k]), k


def run_depth(impl_normals, impl_seeds, impl_planar, impl_cluster, f, name):
    rows, cols = int(f["rows"]), int(f["cols"])
    xyz = f["xyz"]
    nrm = impl_normals(xyz, rows, cols, float(f["min_neighbor_distance"]))
    assert close(nrm, f["normals"]), "normals"
    if "planesupport" in name:
        seeds = impl_seeds(xyz, f["normals"], rows, cols)
        assert np.array_equal(seeds, f["planar_seeds"])
    seeds = f["planar_seeds"]
    lab0 = np.full(rows * cols, -1, np.int32)
    lab1, planes = impl_planar(xyz, f["normals"], rows, cols, seeds, lab0)
    assert np.array_equal(lab1, f["labels_planar"])
    check_regions(planes, f, "planar_", "planar")
    allseeds = np.arange(rows * cols - 1, -1, -1, dtype=np.int32)
    lab2, cl = impl_cluster(xyz, rows, cols, allseeds, lab1, planes.n)
    assert np.array_equal(lab2, f["labels_final"])
    check_regions(cl, f, "cluster_", "cluster")


# ---------------------------------------------------------------- CPU restatement
@pytest.mark.parametrize("name", DEPTH)
def test_oracle_depth(name):
    run_depth(lambda x, r, c, d: oracle.normals_organized(x, r, c, params=oracle.default_normals_params(min_neighbor_distance=d)),
              oracle.seeds_plane_support, oracle.segment_planar,
              lambda x, r, c, s, l, off: oracle.segment_cluster(x, r, c, s, l, id_offset=off), load(name), name)


def test_oracle_lidar():
    f = load("lidar_48x1024_cluster")
    rows, cols = int(f["rows"]), int(f["cols"])
    cfg = oracle.default_cluster_config(squared_distance_threshold=float(f["sqd"]), half_search_window=int(f["hsw"]),
                                        min_region_inliers=int(f["min_inl"]))
    seeds = np.arange(rows * cols - 1, -1, -1, dtype=np.int32)
    lab, cl = oracle.segment_cluster(f["xyz"], rows, cols, seeds, f["labels_in"], cfg=cfg)
    assert np.array_equal(lab, f["labels_out"])
    check_regions(cl, f, "cluster_", "cluster")


def test_oracle_meanshift():
    f = load("meanshift_90x120")
    rows, cols = int(f["rows"]), int(f["cols"])
    lab, regs, _ = oracle.mean_shift(f["xyz"], rows, cols, np.full(rows * cols, -1, np.int32), iterations=int(f["iterations"]))
    assert np.array_equal(lab, f["labels_out"])
    check_regions(regs, f, "ms_", "cluster")
    assert np.all(np.abs(regs.seed_points - f["ms_seed_points"]) <= 1e-4)


# ---------------------------------------------------------------- CUDA path (C ABI)
@pytest.fixture(scope="module")
def ctx():
    import pcseg
    c = pcseg.Context(0)
    yield c
    c.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", DEPTH)
def test_cuda_depth(ctx, name):
    import pcseg
    run_depth(lambda x, r, c, d: ctx.normals_organized(x, r, c, params=dict(min_neighbor_distance=d)),
              ctx.seeds_plane_support, ctx.segment_planar,
              lambda x, r, c, s, l, off: ctx.segment_cluster(x, r, c, s, l, id_offset=off), load(name), name)


@pytest.mark.gpu
def test_cuda_lidar(ctx):
    f = load("lidar_48x1024_cluster")
    rows, cols = int(f["rows"]), int(f["cols"])
    cfg = dict(squared_distance_threshold=float(f["sqd"]), half_search_window=int(f["hsw"]), min_region_inliers=int(f["min_inl"]))
    seeds = np.arange(rows * cols - 1, -1, -1, dtype=np.int32)
    lab, cl = ctx.segment_cluster(f["xyz"], rows, cols, seeds, f["labels_in"], cfg=cfg)
    assert np.array_equal(lab, f["labels_out"])
    check_regions(cl, f, "cluster_", "cluster")


@pytest.mark.gpu
def test_cuda_meanshift(ctx):
    f = load("meanshift_90x120")
    rows, cols = int(f["rows"]), int(f["cols"])
    lab, regs = ctx.mean_shift(f["xyz"], rows, cols, np.full(rows * cols, -1, np.int32), iterations=int(f["iterations"]))
    assert np.array_equal(lab, f["labels_out"])
    check_regions(regs, f, "ms_", "cluster")
    assert np.all(np.abs(regs.seed_points - f["ms_seed_points"]) <= 1e-4)  # 1e-4 of the 1 m bandwidth
