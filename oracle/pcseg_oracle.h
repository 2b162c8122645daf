/*
 * pcseg_oracle.h -- C interface of the CPU ORACLE (test infrastructure only).
 *
 * This is a plain C++17 restatement of the geometric-segmentation hot path of
 * theteamatx/x-edr-pointcloud-segmentation.  It exists only as the checker for
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs.  The product library (libpcseg_b200.so) never links, loads or calls it.
 *
 * Every function cites the reference file:line it follows (paths relative to
 * the reference tree, "ps/" = pointcloud_segmentation/).
 *
 * Parity pin: checked against the reference's own golden vectors
 * (ps/clustering_test.cc, ps/plane_estimator_test.cc, ps/segmentation_test.cc,
 * ps/algorithms_test.cc) in tests/test_oracle_golden.py, and against the
 * reference's own control flow compiled in place against shim headers
 * (oracle/_ref, see oracle/Makefile) in tests/test_oracle_vs_ref.py.
 * The Eigen-3.4.0 arithmetic order (un-vendored dependency) is restated from
 * the published algorithm: bit-level parity with a real Eigen build is UNPINNED.
 */
#ifndef PCSEG_ORACLE_H_
#define PCSEG_ORACLE_H_

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oracle_regions oracle_regions;

typedef struct {
  float min_neighbor_distance;
  float max_neighbor_distance;
  int include_diagonal_neighbors;
  int min_num_support_neighbors;
} oracle_normals_params;

typedef struct {
  float max_plane_distance;
  float min_region_area;
  int min_region_inliers;
  int plane_model_reestimation_period;
  float discontinuity_min_range;
  float discontinuity_max_range;
  float discontinuity_normal_angle_diff;
  float discontinuity_z_diff;
  float discontinuity_z_ratio;
} oracle_planar_config;

typedef struct {
  int min_region_inliers;
  float squared_distance_threshold;
  int half_search_window;
} oracle_cluster_config;

/* 0 = trig through double-evaluate-then-round (default), 1 = native libm
 * float functions (atan2f/cosf/sinf/acosf of this box's glibc). */
void oracle_set_native_libm(int on);

/* ps/algorithms.h:329-375.  normals is in/out: only the range is written. */
int oracle_normals_organized(const float* xyz, int rows, int cols,
                             const double sensor_t[3],
                             const oracle_normals_params* params, int row_begin,
                             int row_end, int col_begin, int col_end,
                             float* normals, float* curvature_or_null);

/* ps/algorithms.h:106-257; returns count, fills out[<=9]. */
int oracle_find_normal_support_neighbors(const float* xyz, int rows, int cols,
                                         int col, int row, float min_d,
                                         float max_d, int diag, int* out);

/* ps/plane_estimator.{h,cc}: stateless replay of an op list on one estimator.
 * ops[i]: 0 = AddPoint(p), 1 = AddPoint(p, w), 2 = SetNormalOrientation(p),
 *         3 = query (appends 10 floats to out: valid, curvature, centroid xyz,
 *             normal xyz, offset, weight-sum),
 *         4 = Clear().
 * args: 4 floats per op (x, y, z, w).  Returns number of queries written. */
int oracle_plane_estimator_replay(const int* ops, const float* args, int n_ops,
                                  float* out, int out_cap);

/* Same with the rest of the class (ps/plane_estimator.cc:128-182): four estimator slots and
 *         5 = select active slot x, 6 = Merge(slot x), 7 = Transform(pose x),
 *         8 = Merge(slot x, pose y).  poses: 7 doubles each (quaternion xyzw, translation),
 * an eigenmath::Pose3d.  Returns -3 where the reference would CHECK-abort (zero weight). */
int oracle_plane_estimator_replay2(const int* ops, const float* args, int n_ops,
                                   const double* poses, int n_poses,
                                   float* out, int out_cap);

/* Eigen 3.4.0 SelfAdjointEigenSolver<Matrix3f>::computeDirect restated.
 * a = 6 floats (m00,m10,m20,m11,m21,m22); evals[3]; evecs col-major 9. */
void oracle_eig3(const float* a, float* evals, float* evecs);

/* ps/algorithms.h:264-292.  indices may be NULL (NoIndices). */
float oracle_polygon_area(const float* xyz, int n_points, const int* indices,
                          int n_indices, float* cumulative_normal_or_null);

/* ps/segmentation.cc:21-26 */
void oracle_prepare_unlabeled_labels(int* labels, int n);

/* ps/segmentation.h:239-292 with RegionT = PlanarRegion (ps/planar_region.h). */
int oracle_segment_planar(const float* xyz, const float* normals, int rows,
                          int cols, const float robot_q_xyzw[4],
                          const int* seeds, int n_seeds,
                          const oracle_planar_config* cfg, int id_offset,
                          int* labels, oracle_regions** out);

/* diagnostics: growth attempts / queue pops by inlier-count bucket since the last reset */
void oracle_debug_attempt_hist(long long out[16], int reset);

/* ps/segmentation.h:239-292 with RegionT = ClusterRegion (ps/cluster_region.h). */
int oracle_segment_cluster(const float* xyz, int rows, int cols,
                           const int* seeds, int n_seeds,
                           const oracle_cluster_config* cfg, int id_offset,
                           int* labels, oracle_regions** out);

/* Fixed-radius Euclidean clustering of an unorganized cloud through a voxel grid (extension with
 * library-defined semantics, csrc/voxel_cluster.cuh): connected components of "squared distance <
 * radius^2" over the unlabeled finite points; regions numbered by smallest member index. */
int oracle_cluster_unorganized(const float* xyz, int n, float radius,
                               int min_region_inliers, int id_offset,
                               int* labels, oracle_regions** out);

/* ps/mean_shift_segmentation.h:207-330.  modes_out (nullable): per seed
 * 8 floats {mode xyz, index row, index col, intensity, valid, cloud index},
 * n_modes = number of seed points (unlabeled & non-NaN), in seed order.
 * sort_mode: 0 = std::sort (reference), 1 = std::stable_sort (patch P2). */
int oracle_mean_shift(const float* xyz, int rows, int cols,
                      const oracle_cluster_config* cfg, int iterations,
                      int id_offset, int sort_mode, int* labels,
                      oracle_regions** out, float* modes_out, int modes_cap,
                      int* n_modes);

/* ps/segmentation.h:190-230 (quirk Q1 reproduced literally). */
int oracle_seeds_plane_support(const float* xyz, const float* normals, int rows,
                               int cols, int neighborhood_size,
                               float max_plane_distance, int min_support,
                               int* seeds_out, int cap);
/* ps/segmentation.h:136-184 (quirk Q1 reproduced literally). */
int oracle_seeds_average_normals(const float* xyz, const float* normals,
                                 int rows, int cols, int neighborhood_size,
                                 int min_num_valid_normals,
                                 float min_avg_normal_length, int* seeds_out,
                                 int cap);

/* The std::sort call of ps/mean_shift_segmentation.h:276-280 on its own (port library only):
 * order_out[k] = input position of the element sorted to place k; sort_mode as above. */
void oracle_sort_order(const double* intensities, int n, int sort_mode, int* order_out);

/* ps/planar_region.h:478-519 FindSeedPointsFromLastPlanarRegions; previous regions given by
 * inlier count, centroid and plane normal; pose = current_pose_prev (Pose3f).  seeds_out has room
 * for n_prev entries; returns the number appended. */
int oracle_seeds_last_planar_regions(const float* xyz, const float* normals, int n_points, int n_prev,
                                     const int* prev_inlier_counts, const float* prev_centroids,
                                     const float* prev_normals, const float q_xyzw[4], const float t[3],
                                     float max_distance, float max_normal_difference_angle, int* seeds_out);

/* ps/multichannel_cloud.cc:53-68 MultichannelCloud::TransformInPlace (either array nullable).
 * Port only: the reference's class needs the generated proto headers. */
void oracle_transform_in_place(float* points, float* normals, int n,
                               const double q_xyzw[4], const double t[3]);

/* Result accessors (pointers valid until oracle_regions_free). */
int oracle_regions_count(const oracle_regions*);
const int* oracle_regions_inlier_offsets(const oracle_regions*);   /* n+1 */
const int* oracle_regions_inlier_indices(const oracle_regions*);
const int* oracle_regions_label_ids(const oracle_regions*);        /* n */
const int* oracle_regions_seed_indices(const oracle_regions*);     /* n (planar) */
const float* oracle_regions_planes(const oracle_regions*);         /* 4n */
const float* oracle_regions_centroids(const oracle_regions*);      /* 3n */
const float* oracle_regions_areas(const oracle_regions*);          /* n */
const float* oracle_regions_estimators(const oracle_regions*);     /* 10n */
const int* oracle_regions_boundary_offsets(const oracle_regions*); /* n+1 */
const int* oracle_regions_boundary_indices(const oracle_regions*);
const int* oracle_regions_discontinuous_offsets(const oracle_regions*);
const int* oracle_regions_discontinuous_indices(const oracle_regions*); /* sorted */
const int* oracle_regions_hull_offsets(const oracle_regions*);     /* n+1, in points */
const float* oracle_regions_hull_points(const oracle_regions*);    /* xyz */
const float* oracle_regions_seed_points(const oracle_regions*);    /* 3n (mean shift) */
const float* oracle_regions_bboxes(const oracle_regions*);         /* 6n (unorganized clustering) */
void oracle_regions_free(oracle_regions*);

#ifdef __cplusplus
}
#endif
#endif  /* PCSEG_ORACLE_H_ */
