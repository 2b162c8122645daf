/*
 * pcseg_b200.h -- C ABI of the B200-native geometric segmentation hot path.
 *
 * Drop-in boundary for the hot path of theteamatx/x-edr-pointcloud-segmentation
 * ("ps/" = pointcloud_segmentation/ in the reference tree).  Plain pointers and
 * sizes only.  Clouds use the reference's boundary layout (ps/cloud.h:37-41,74;
 * ps/multichannel_cloud_test.cc:209-218): column-major grid,
 * index = col * rows + row, AoS float32 xyz (12 B per point), NaN = invalid,
 * int32 labels with the sentinels of ps/segmentation.h:36-45.
 *
 * Every data pointer may be a HOST pointer or a DEVICE pointer on the context's
 * GPU (detected with cudaPointerGetAttributes); outputs are written to wherever
 * the caller's pointer lives.  All entry points return 0 on success; a nonzero
 * status names the reference CHECK that would have aborted the process
 * (ps/planar_region.h:257-268, ps/cluster_region.h:152-159,
 * ps/mean_shift_segmentation.h:212-215) or a CUDA / capacity error.  There is
 * no CPU fallback: without a usable CUDA device pcseg_create fails.
 */
#ifndef PCSEG_B200_H_
#define PCSEG_B200_H_

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pcseg_ctx pcseg_ctx;
typedef struct pcseg_regions pcseg_regions;
typedef struct pcseg_batch pcseg_batch;

enum {
  PCSEG_OK = 0,
  PCSEG_ERR_NULL = 1,        /* CHECK(ptr) */
  PCSEG_ERR_DIMS = 2,        /* CHECK_EQ on rows/cols, CHECK_GT(rows, 1) */
  PCSEG_ERR_CONFIG = 3,      /* config value the reference turns into UB (quirk Q8) */
  PCSEG_ERR_CUDA = 4,        /* CUDA runtime error (see pcseg_last_error) */
  PCSEG_ERR_CAPACITY = 5,    /* an internal per-frame buffer overflowed */
  PCSEG_ERR_UNSUPPORTED = 6, /* documented limit of this implementation */
  PCSEG_ERR_SEED_RANGE = 7   /* seed index outside [0, rows*cols) */
};

/* ps/algorithms.h:313-322 ComputeNormalsParams */
typedef struct {
  float min_neighbor_distance;      /* 0.1 */
  float max_neighbor_distance;      /* 1.0 */
  int include_diagonal_neighbors;   /* 1   */
  int min_num_support_neighbors;    /* 4   */
} pcseg_normals_params;

/* ps/region_segmentation_config.proto:42-72 PlanarRegionConfigProto with the
 * defaults of ps/planar_region.h:93-121 already applied by the caller. */
typedef struct {
  float max_plane_distance;               /* 0.05 */
  float min_region_area;                  /* 0.05 */
  int min_region_inliers;                 /* 5    */
  int plane_model_reestimation_period;    /* 30   */
  float discontinuity_min_range;          /* 1.2  */
  float discontinuity_max_range;          /* 4.0  */
  float discontinuity_normal_angle_diff;  /* 5.0  */
  float discontinuity_z_diff;             /* 0.05 */
  float discontinuity_z_ratio;            /* 0.7  */
} pcseg_planar_config;

/* ps/region_segmentation_config.proto:22-39 ClusterRegionConfigProto with the
 * defaults of ps/cluster_region.h:53-63 applied. */
typedef struct {
  int min_region_inliers;             /* 7   */
  float squared_distance_threshold;   /* 1.0 */
  int half_search_window;             /* 1   (<= 7 supported) */
} pcseg_cluster_config;

void pcseg_default_normals_params(pcseg_normals_params*);
void pcseg_default_planar_config(pcseg_planar_config*);
void pcseg_default_cluster_config(pcseg_cluster_config*);

/* One context per GPU; calls on one context are serialised on its stream. */
int pcseg_create(int device, pcseg_ctx** out);
void pcseg_destroy(pcseg_ctx*);
const char* pcseg_status_string(int status);
const char* pcseg_last_error(const pcseg_ctx*);
/* Number of kernels this context has launched so far (bench.py gpu_launches). */
long long pcseg_launch_count(const pcseg_ctx*);
/* The CUDA stream (cudaStream_t as void*) the context launches on. */
void* pcseg_stream(const pcseg_ctx*);

/* Replaces ps/segmentation.cc:21-26 PrepareUnlabeledLabels: fill with -1. */
int pcseg_prepare_unlabeled_labels(pcseg_ctx*, int* labels, int n);

/* Replaces ps/algorithms.h:329-375 ComputeNormalsOrganized (both overloads;
 * pass the full range for the second).  Only the translation of
 * point_cloud_pose_sensor is used by the reference (:355).  normals is in/out:
 * only cells inside [row_begin,row_end) x [col_begin,col_end) are written.
 * curvature (nullable) receives PlaneEstimator::Curvature() per point
 * (ps/plane_estimator.cc:106-111; extension output, NaN where the normal is). */
int pcseg_normals_organized(pcseg_ctx*, const float* xyz, int rows, int cols,
                            const double sensor_translation[3],
                            const pcseg_normals_params* params, int row_begin,
                            int row_end, int col_begin, int col_end,
                            float* normals, float* curvature);

/* Replaces ps/segmentation.h:239-292 SegmentRegions<PlanarRegion<>> with
 * ps/planar_region.h:123-268 as the region model.  robot_q_xyzw is the
 * quaternion of InputData::robot_pose_point_cloud (only used for the
 * discontinuity side output, ps/planar_region.h:390).  labels is in/out.
 * Seeds are consumed back to front (ps/segmentation.h:254-256). */
int pcseg_segment_planar(pcseg_ctx*, const float* xyz, const float* normals,
                         int rows, int cols, const float robot_q_xyzw[4],
                         const int* seeds, int n_seeds,
                         const pcseg_planar_config* cfg, int initial_id_offset,
                         int* labels, pcseg_regions** out);

/* Replaces SegmentRegions<ClusterRegion<>> (ps/cluster_region.h:47-176). */
int pcseg_segment_cluster(pcseg_ctx*, const float* xyz, int rows, int cols,
                          const int* seeds, int n_seeds,
                          const pcseg_cluster_config* cfg, int initial_id_offset,
                          int* labels, pcseg_regions** out);

/* Replaces the iteration phase of ps/mean_shift_segmentation.h:217-260.
 * modes_out: 8 floats per cloud cell {mode xyz, index row, index col,
 * intensity, valid, is_seed}; cells that are not seeds have is_seed = 0. */
int pcseg_mean_shift_modes(pcseg_ctx*, const float* xyz, int rows, int cols,
                           const int* labels, int iterations, float* modes_out);

/* Same for the seed cells [cell_begin, cell_end) only (contiguous linear-index
 * shard; seeds are mutually independent while labels are read-only,
 * ps/mean_shift_segmentation.h:232-260).  modes_out holds
 * (cell_end - cell_begin) records.  Used to shard one large cloud's seeds over
 * several GPUs (cloud + labels replicated, records all-gathered by the caller). */
int pcseg_mean_shift_modes_range(pcseg_ctx*, const float* xyz, int rows, int cols,
                                 const int* labels, int iterations, int cell_begin,
                                 int cell_end, float* modes_out);

/* Sort + growth phase (ps/mean_shift_segmentation.h:262-329) on mode records
 * produced by pcseg_mean_shift_modes[_range] (rows*cols records, e.g. after an
 * all-gather of the shards).  labels is in/out. */
int pcseg_mean_shift_from_modes(pcseg_ctx*, const float* xyz, int rows, int cols,
                                const float* modes,
                                const pcseg_cluster_config* cfg,
                                int initial_region_id_offset, int sort_mode,
                                int* labels, pcseg_regions** out);

/* The sort of ps/mean_shift_segmentation.h:262-280 on its own: intensities[i] >= 0 marks element i
 * as a valid seed (window counts: integers <= 255); order_out receives the positions of the valid
 * seeds in sorted order, *n_out their number.  sort_mode 0 = the permutation of the reference's
 * std::sort call for libstdc++ (introsort replayed on the device); 1 = stable.
 * PCSEG_ERR_UNSUPPORTED if std::sort would enter its heapsort fall-back (depth limit 2*log2 n). */
int pcseg_mean_shift_sort_order(pcseg_ctx*, const float* intensities, int n, int sort_mode,
                                int* order_out, int* n_out);

/* Replaces ps/mean_shift_segmentation.h:207-330 SlidingMeanShift (iteration
 * phase + sort + growth).  sort_mode 0 = the reference's std::sort permutation
 * (libstdc++ introsort replayed on the device), 1 = stable.  Everything runs on the device.
 * Label values >= 0x40000000 in the caller's grid are refused (PCSEG_ERR_UNSUPPORTED): the growth
 * phase uses that range for temporary labels (environment PCSEG_MS_GROWTH=serial lifts the limit). */
int pcseg_mean_shift(pcseg_ctx*, const float* xyz, int rows, int cols,
                     const pcseg_cluster_config* cfg, int iterations,
                     int initial_region_id_offset, int sort_mode, int* labels,
                     pcseg_regions** out);

/* Replaces ps/segmentation.h:190-230 FindSeedPointsFromPlaneSupport (quirk Q1
 * reproduced literally).  Returns the number of seeds in *n_out. */
int pcseg_seeds_plane_support(pcseg_ctx*, const float* xyz, const float* normals,
                              int rows, int cols, int neighborhood_size,
                              float max_plane_distance, int min_num_support,
                              int* seeds_out, int cap, int* n_out);

/* Replaces ps/segmentation.h:137-184 FindSeedPointsFromAverageNormals (with the sliding window
 * of :79-120; quirk Q1 reproduced literally, so rows <= cols is required as the reference's
 * transposed reads would leave the buffer otherwise).  Seeds come out in the reference's order
 * (row outer, col inner).  Returns the number of seeds in *n_out (may exceed cap; only cap are
 * written).  neighborhood_size <= 32. */
int pcseg_seeds_average_normals(pcseg_ctx*, const float* xyz, const float* normals,
                                int rows, int cols, int neighborhood_size,
                                int min_num_valid_normals, float min_avg_normal_length,
                                int* seeds_out, int cap, int* n_out);

/* No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The same computation and the same seeds as pcseg_seeds_average_normals, but the entry point
 * allocates NOTHING (no cudaMalloc / cudaFree / allocating library routine): its scratch (one
 * flag byte per window position) lives in a file-scope `static __device__` array of
 * PCSEG_NOALLOC_MAX_CELLS bytes that every GPU loading the library holds until exit.
 *   - normals_dev, seeds_out_dev, n_out_dev are DEVICE pointers on the context's device (host
 *     pointers would need a staging allocation: PCSEG_ERR_UNSUPPORTED).  seeds_out_dev holds cap
 *     ints and cap must cover every window position, cap >= (rows - 2*(neighborhood_size/2)) * cols
 *     (PCSEG_ERR_CAPACITY otherwise); *n_out_dev receives the number of seeds.
 *   - rows * cols > PCSEG_NOALLOC_MAX_CELLS -> PCSEG_ERR_CAPACITY with a message in pcseg_last_error.
 *   - The call returns after the context's stream has drained (cudaStreamSynchronize).
 *   - ALL calls (any context, any thread) on one device share the array, which keeps the last
 *     call's contents; every flag the second kernel reads is written by the first kernel of the same
 *     call, so nothing has to start at zero.  Calls that overlap on one GPU race: serialise them. */
#define PCSEG_NOALLOC_MAX_CELLS (1 << 20)
int pcseg_seeds_average_normals_noalloc(pcseg_ctx*, const float* normals_dev, int rows, int cols,
                                        int neighborhood_size, int min_num_valid_normals,
                                        float min_avg_normal_length, int* seeds_out_dev, int cap,
                                        int* n_out_dev);

/* Replaces ps/planar_region.h:478-519 FindSeedPointsFromLastPlanarRegions.  The previous
 * regions are given by what the function reads of them: inlier_indices().size(), centroid(),
 * plane().normal().  current_pose_prev = (q_xyzw, translation), an eigenmath::Pose3f.
 * The seeds found are APPENDED to seeds_inout[*n_inout ...] (the reference does not clear its
 * output vector) in the reference's order: previous regions by ascending inlier count,
 * insertion order within a count (std::multimap). */
int pcseg_seeds_last_planar_regions(pcseg_ctx*, const float* xyz, const float* normals,
                                    int n_points, int n_prev, const int* prev_inlier_counts,
                                    const float* prev_centroids, const float* prev_plane_normals,
                                    const float q_xyzw[4], const float translation[3],
                                    float max_distance, float max_normal_difference_angle,
                                    int* seeds_inout, int cap, int* n_inout);

/* ---- PlaneEstimator (ps/plane_estimator.h:37-120, ps/plane_estimator.cc:46-229) ----
 * Query row used below: {plane valid, curvature, centroid xyz, plane normal xyz, plane offset,
 * cumulative weight}.  Accumulator state: {covariance accumulator[6], cumulative centroid xyz,
 * cumulative weight} -- the same 10 floats as pcseg_regions_estimators and
 * PlaneEstimatorProto (ps/plane_estimator.proto:22-32). */

/* Least-squares plane fits of n_regions index sets at once (north_star item 2): per region
 * Clear(); [SetNormalOrientation(hint);] AddPoints(points, indices) in list order -- sequential
 * fp32 sums, as the reference (ps/plane_estimator.h:97-103) -- then Plane()/Centroid()/Curvature().
 * weights (nullable) per index entry: weighted AddPoint (ps/plane_estimator.cc:67-77).
 * hints (nullable) 3 floats per region.  query_out: 10 floats per region; state_out (nullable)
 * 10 floats per region.  A region with no entries returns the state right after Clear(). */
int pcseg_fit_planes(pcseg_ctx*, const float* xyz, int n_points, const int* offsets,
                     const int* indices, int n_regions, const float* weights,
                     const float* hints, float* query_out, float* state_out);

/* The whole class, op by op on ONE estimator chain (the reference's semantics including the
 * lazy recomputation and the kept normal orientation): ops[i] with 4 float args each:
 *   0 AddPoint(xyz)  1 AddPoint(xyz, w)  2 SetNormalOrientation(xyz)  3 query (appends a row)
 *   4 Clear()  5 select active estimator slot x (4 slots)  6 Merge(slot x)  7 Transform(pose x)
 *   8 Merge(slot x, pose y).   poses: 7 doubles each (quaternion xyzw, translation): Pose3d.
 * *n_queries receives the number of rows written; PCSEG_ERR_CONFIG where the reference
 * CHECK-aborts (Plane()/Transform() with zero weight), PCSEG_ERR_CAPACITY if cap rows are not enough. */
int pcseg_plane_estimator_replay(pcseg_ctx*, const int* ops, const float* args, int n_ops,
                                 const double* poses, int n_poses, float* query_out, int cap,
                                 int* n_queries);

/* Transform (ps/plane_estimator.cc:142-182) of n accumulator states in place; poses: 7 doubles,
 * one per state (pose_per_state != 0) or one shared.  PCSEG_ERR_CONFIG if a state has weight <= 0. */
int pcseg_plane_estimators_transform(pcseg_ctx*, float* states, int n, const double* poses,
                                     int pose_per_state);
/* Merge (ps/plane_estimator.cc:128-140): dst[i] += src[i], src[i] first transformed by its pose
 * when poses is non-null (this_pose_other). */
int pcseg_plane_estimators_merge(pcseg_ctx*, float* dst_states, const float* src_states, int n,
                                 const double* poses, int pose_per_state);

/* ---- Unorganized clouds (extension; SURVEY.md 8(f) rank 4, BASELINE.json config 4): fixed-radius
 * Euclidean clustering through a voxel-hash grid (voxel = radius).  The reference has no
 * unorganized path (its ClusterRegion needs a grid, ps/cluster_region.h:122-150), so the semantics
 * are defined by this library (csrc/voxel_cluster.cuh) and restated by the CPU oracle:
 * participating points = labels == kUnlabeled and finite; neighbours = squared distance < radius^2
 * (fp32, the reduction order of ps/cluster_region.h:94); clusters = connected components with
 * >= min_region_inliers points; ids = initial_id_offset + rank by smallest member index; inlier
 * lists ascending.  Result accessors: count, inlier_offsets / inlier_indices, label_ids, centroids
 * (sum of the member points / count) and bboxes (min xyz, max xyz) -- north_star item (4). ---- */
typedef struct {
  float radius;
  int min_region_inliers;
} pcseg_voxel_cluster_config;
int pcseg_cluster_unorganized(pcseg_ctx*, const float* xyz, int n_points,
                              const pcseg_voxel_cluster_config* cfg, int initial_id_offset,
                              int* labels, pcseg_regions** out);

/* ---- Batched pipeline (extension: the reference has no batch entry point;
 * a caller looping ComputeNormalsOrganized -> SegmentRegions<PlanarRegion> ->
 * SegmentRegions<ClusterRegion> over independent frames binds this instead).
 * Frames are contiguous: xyz[f][rows*cols][3]. ---- */
enum {
  PCSEG_SEEDS_ALL_DESCENDING = 0, /* seeds[k] = N-1-k, as ps/clustering_test.cc:57-59 */
  PCSEG_SEEDS_PLANE_SUPPORT = 1,  /* FindSeedPointsFromPlaneSupport defaults, on device */
  PCSEG_SEEDS_EXPLICIT = 2        /* caller-provided list shared by all frames */
};
typedef struct {
  pcseg_normals_params normals;
  pcseg_planar_config planar;
  pcseg_cluster_config cluster;
  double sensor_translation[3];
  float robot_q_xyzw[4];
  int planar_seed_mode;
  const int* planar_seeds; /* PCSEG_SEEDS_EXPLICIT only */
  int n_planar_seeds;
  int run_cluster_stage;   /* seeds all descending, id offset = #planes of the frame */
} pcseg_pipeline_params;

int pcseg_batch_create(pcseg_ctx*, int n_frames, int rows, int cols,
                       pcseg_batch** out);
void pcseg_batch_destroy(pcseg_batch*);
/* normals_out / labels_out nullable (results then stay on the device only). */
int pcseg_batch_run(pcseg_batch*, const float* xyz,
                    const pcseg_pipeline_params* params, float* normals_out,
                    int* labels_out);
/* Asynchronous form of pcseg_batch_run: submit enqueues the whole pipeline (copies included)
 * on the workspace's own streams and returns; wait blocks until its results are complete.
 * Two workspaces of one context used alternately overlap the host copies of one batch with
 * the kernels of the other (a caller's frame stream: bench.py's e2e arm).  The buffers passed
 * to submit must stay valid (and pinned, for real overlap) until wait returns; submitting to
 * a workspace that is still in flight waits for it first. */
int pcseg_batch_submit(pcseg_batch*, const float* xyz,
                       const pcseg_pipeline_params* params, float* normals_out,
                       int* labels_out);
int pcseg_batch_wait(pcseg_batch*);
/* The CUDA stream (cudaStream_t as void*) the workspace's kernels are launched on. */
void* pcseg_batch_stream(const pcseg_batch*);
/* Per-frame counters after a run: {n_planar, n_cluster, planar_inliers,
 * cluster_inliers, boundary_len, error_flags, grow_steps, refits}. */
int pcseg_batch_frame_stats(pcseg_batch*, int* stats8_per_frame);
/* Device time of the last run's stages in ms (CUDA events on the context's
 * stream): {normals + fix-up, label init + seed finder, planar growth,
 * cluster growth}. */
int pcseg_batch_stage_times(pcseg_batch*, float* ms4);
/* Diagnostics (only in -DPCSEG_PROFILE builds, else PCSEG_ERR_UNSUPPORTED):
 * clock64 totals / event counts of the planar growth roles, 16 per frame. */
int pcseg_batch_profile(pcseg_batch*, long long* out16_per_frame);
/* Materialises one frame's regions (kind 0 = planar, 1 = cluster). */
int pcseg_batch_regions(pcseg_batch*, int frame, int kind, pcseg_regions** out);
/* Device pointers of the batch's resident results (labels int32[F*N],
 * normals float[F*N*3]); valid until pcseg_batch_destroy. */
int* pcseg_batch_device_labels(pcseg_batch*);
float* pcseg_batch_device_normals(pcseg_batch*);

/* ---- Batched region results: everything a PlanarRegion / ClusterRegion holds
 * (ps/planar_region.h:427-450, ps/cluster_region.h:161-168) for ALL frames of a
 * batch, packed on the device (one pass) and copied with one transfer per array.
 * With want_inlier_xyz / want_boundary_xyz the gathers of
 * CreatePlanarDetectedObjectProto (ps/detected_objects.cc:21-48: inlier xyz in
 * inlier order, discontinuous boundary index -> position in that list) and
 * CreateClusterDetectedObjectProto (ps/detected_objects.h:62-74) run on the
 * device too.  Region r of the batch: frame-major, region order within a frame =
 * the order SegmentRegions pushed them (ps/segmentation.h:278). ---- */
typedef struct {
  int frame, label_id, seed_index;
  int inlier_begin, inlier_count;        /* into planar_inlier_indices / planar_inlier_xyz */
  int boundary_begin, boundary_count;    /* into boundary_indices / boundary_xyz / boundary_discontinuous */
  int discontinuous_begin, discontinuous_count; /* into discontinuous_indices / discontinuous_positions */
  int hull_begin, hull_count;            /* into hull_points (projected_boundary_points_) */
  float plane[4];                        /* normal xyz, offset */
  float centroid[3];
  float area;
  float estimator[10];                   /* PlaneEstimator: 6 second moments, sum xyz, weight */
} pcseg_planar_record;

typedef struct {
  int frame, label_id;
  int inlier_begin, inlier_count;        /* into cluster_inlier_indices / cluster_inlier_xyz */
  /* extension outputs (no reference field): centroid = sum of the inlier points / count, lane-strided
   * double sums; axis-aligned extent */
  float centroid[3];
  float bbox_min[3], bbox_max[3];
} pcseg_cluster_record;

typedef struct {
  /* in: which optional gathers to run */
  int want_inlier_xyz;    /* planar_inlier_xyz, cluster_inlier_xyz */
  int want_boundary_xyz;  /* boundary_xyz */
  int want_hull;          /* hull_points */
  int want_cluster_stats; /* centroid / bbox of pcseg_cluster_record (else zero) */
  /* in: caller buffers (host -- pinned for overlap -- or device) and their capacities in
   * ELEMENTS (records, indices, points).  A null buffer skips that array. */
  int* frame_stats;                      /* [F][8], as pcseg_batch_frame_stats */
  int* frame_offsets;                    /* [5][F+1] exclusive prefix sums over frames of: planar regions,
                                            planar inliers, boundary entries, cluster regions, cluster inliers */
  pcseg_planar_record* planar;           int planar_cap;
  int* planar_inlier_indices;            int planar_inlier_cap;
  float* planar_inlier_xyz;              /* 3 per inlier, same capacity */
  int* boundary_indices;                 int boundary_cap;
  float* boundary_xyz;                   /* 3 per boundary entry */
  unsigned char* boundary_discontinuous; /* flag per boundary entry (ps/planar_region.h:360-417) */
  int* discontinuous_indices;            /* cloud index, ascending position in the inlier list; capacity boundary_cap;
                                            a region's entries start at discontinuous_begin (= its boundary_begin) */
  int* discontinuous_positions;          /* position in the region's inlier list (ps/detected_objects.cc:41-45) */
  float* hull_points;                    /* 3 per hull vertex; capacity boundary_cap; hull_begin = boundary_begin */
  pcseg_cluster_record* cluster;         int cluster_cap;
  int* cluster_inlier_indices;           int cluster_inlier_cap;
  float* cluster_inlier_xyz;
  /* out (valid after pcseg_batch_wait): totals over the batch */
  int n_planar, n_planar_inliers, n_boundary, n_cluster, n_cluster_inliers;
  int error_flags;                       /* OR over frames of FrameStats.error_flags */
} pcseg_batch_results;

/* pcseg_batch_submit + region results: `results` (and every buffer it names) must stay valid until
 * pcseg_batch_wait returns; wait fills the totals and returns PCSEG_ERR_CAPACITY when a caller
 * buffer was too small (totals are still filled in, nothing is written past a capacity). */
int pcseg_batch_submit_regions(pcseg_batch*, const float* xyz,
                               const pcseg_pipeline_params* params, float* normals_out,
                               int* labels_out, pcseg_batch_results* results);

/* Replaces MultichannelCloud::TransformInPlace (ps/multichannel_cloud.cc:53-68): points <- pose * point,
 * normals <- rotationMatrix(pose) * normal, with new_pose_current = (quaternion xyzw, translation) given
 * in double and cast to float first (Pose3d::cast<float>, :55).  Either array may be null. */
int pcseg_transform_in_place(pcseg_ctx*, float* points_xyz, float* normals_xyz, int n,
                             const double q_xyzw[4], const double translation[3]);

/* The gather of CreateClusterDetectedObjectProto (ps/detected_objects.h:62-74) /
 * CreatePlanarDetectedObjectProto (ps/detected_objects.cc:36-46) for one index list:
 * out_xyz[k] = xyz[indices[k]]; discontinuous (nullable, sorted or not) ->
 * positions_out[j] ascending = positions k whose index is in the set; *n_positions = count. */
int pcseg_gather_points(pcseg_ctx*, const float* xyz, int n_points, const int* indices, int n_indices,
                        float* out_xyz, const int* discontinuous, int n_discontinuous,
                        int* positions_out, int* n_positions);

/* ---- Result accessors (host pointers valid until pcseg_regions_free).
 * Mirror PlanarRegion's accessors (ps/planar_region.h:427-450) and
 * ClusterRegion's (ps/cluster_region.h:161-168) as CSR arrays. ---- */
int pcseg_regions_count(const pcseg_regions*);
const int* pcseg_regions_inlier_offsets(const pcseg_regions*);        /* n+1 */
const int* pcseg_regions_inlier_indices(const pcseg_regions*);
const int* pcseg_regions_label_ids(const pcseg_regions*);             /* n   */
const int* pcseg_regions_seed_indices(const pcseg_regions*);          /* n   */
const float* pcseg_regions_planes(const pcseg_regions*);              /* 4n: normal, offset */
const float* pcseg_regions_centroids(const pcseg_regions*);           /* 3n  */
const float* pcseg_regions_areas(const pcseg_regions*);               /* n   */
const float* pcseg_regions_estimators(const pcseg_regions*);          /* 10n: acc[6], sum xyz, weight */
const int* pcseg_regions_boundary_offsets(const pcseg_regions*);      /* n+1 */
const int* pcseg_regions_boundary_indices(const pcseg_regions*);
const int* pcseg_regions_discontinuous_offsets(const pcseg_regions*); /* n+1 */
const int* pcseg_regions_discontinuous_indices(const pcseg_regions*); /* a set, listed in inlier order */
const int* pcseg_regions_hull_offsets(const pcseg_regions*);          /* n+1 (points) */
const float* pcseg_regions_hull_points(const pcseg_regions*);         /* xyz */
const float* pcseg_regions_seed_points(const pcseg_regions*);         /* 3n (mean shift) */
const float* pcseg_regions_bboxes(const pcseg_regions*);              /* 6n (pcseg_cluster_unorganized) */
void pcseg_regions_free(pcseg_regions*);

#ifdef __cplusplus
}
#endif
#endif /* PCSEG_B200_H_ */
