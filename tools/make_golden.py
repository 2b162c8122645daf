#!/usr/bin/env python
"""Generates tests/golden/*.npz: golden input/output vectors produced by the
REFERENCE's own code (oracle/_ref/libref.so = the reference's .h/.cc compiled in
place from /root/reference against oracle/shims, patch P1 only).

Run in the build container (needs /root/reference):  make -C oracle ref && python tools/make_golden.py
The fixtures are committed; tests read them without the reference tree:
  tests/test_golden_fixtures.py  (-m "not gpu": restatement vs fixtures; -m gpu: CUDA vs fixtures)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "oracle"), os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200")]
import oracle  # noqa: E402
import synth   # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def region_arrays(prefix, r, kind):
    d = {prefix + "inlier_offsets": r.inlier_offsets, prefix + "inlier_indices": r.inlier_indices,
         prefix + "label_ids": r.label_ids}
    if kind == "planar":
        for k in ("seed_indices", "planes", "centroids", "areas", "estimators", "boundary_offsets",
                  "boundary_indices", "disc_offsets", "disc_indices", "hull_offsets", "hull_points"):
            d[prefix + k] = getattr(r, k)
    if kind == "meanshift":
        d[prefix + "seed_points"] = r.seed_points
    return d


def depth_case(name, rows, cols, seed, min_d, seed_mode):
    xyz = synth.depth_frame(rows, cols, seed=seed)
    nparams = oracle.default_normals_params(min_neighbor_distance=min_d)
    with oracle.reference():
        normals = oracle.normals_organized(xyz, rows, cols, params=nparams)
        if seed_mode == "planesupport":
            seeds = oracle.seeds_plane_support(xyz, normals, rows, cols)
        else:
            seeds = np.arange(rows * cols - 1, -1, -1, dtype=np.int32)
        lab0 = np.full(rows * cols, -1, np.int32)
        lab1, planes = oracle.segment_planar(xyz, normals, rows, cols, seeds, lab0)
        allseeds = np.arange(rows * cols - 1, -1, -1, dtype=np.int32)
        lab2, clusters = oracle.segment_cluster(xyz, rows, cols, allseeds, lab1, id_offset=planes.n)
    d = dict(xyz=xyz, rows=rows, cols=cols, min_neighbor_distance=np.float32(min_d), normals=normals,
             planar_seeds=seeds, labels_planar=lab1, labels_final=lab2)
    d.update(region_arrays("planar_", planes, "planar"))
    d.update(region_arrays("cluster_", clusters, "cluster"))
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
    print(name, "planes", planes.n, "clusters", clusters.n, "nan normals %.3f" % np.isnan(normals[:, 0]).mean())


def lidar_case(name, rows, cols):
    xyz, labels = synth.c4(rows, cols, seed=3234, n_objects=120)
    cfg = oracle.default_cluster_config(squared_distance_threshold=0.25, half_search_window=2, min_region_inliers=10)
    seeds = np.arange(rows * cols - 1, -1, -1, dtype=np.int32)
    with oracle.reference():
        lab, cl = oracle.segment_cluster(xyz, rows, cols, seeds, labels, cfg=cfg, id_offset=0)
    d = dict(xyz=xyz, rows=rows, cols=cols, labels_in=labels, labels_out=lab, sqd=np.float32(0.25), hsw=2, min_inl=10)
    d.update(region_arrays("cluster_", cl, "cluster"))
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
    print(name, "clusters", cl.n)


def meanshift_case(name, rows, cols):
    xyz = synth.c5(rows, cols, seed=4234, n_blobs=8, pitch=0.12)
    lab0 = np.full(rows * cols, -1, np.int32)
    with oracle.reference():
        lab, regs, _ = oracle.mean_shift(xyz, rows, cols, lab0, iterations=5)
    d = dict(xyz=xyz, rows=rows, cols=cols, iterations=5, labels_out=lab)
    d.update(region_arrays("ms_", regs, "meanshift"))
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
    print(name, "regions", regs.n, "labelled", int((lab >= 0).sum()))


def main():
    if not oracle.ref_available():
        sys.exit("oracle/_ref/libref.so missing: run `make -C oracle ref` where /root/reference exists")
    os.makedirs(OUT, exist_ok=True)
    depth_case("depth_96x128_alldesc", 96, 128, 1234, 0.05, "alldesc")
    depth_case("depth_96x128_planesupport", 96, 128, 1300, 0.05, "planesupport")
    lidar_case("lidar_48x1024_cluster", 48, 1024)
    meanshift_case("meanshift_90x120", 90, 120)


if __name__ == "__main__":
    main()
