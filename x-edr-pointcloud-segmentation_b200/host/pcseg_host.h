// pcseg_host.h -- C++17 host-side mirror of the reference's segmentation API for
// the hot path, on top of the C ABI (include/pcseg_b200.h).
//
// Same names, argument meaning and error behaviour as the reference (precondition
// failure = process abort like absl CHECK,
// ps/planar_region.h:257-268, ps/cluster_region.h:152-159,
// ps/mean_shift_segmentation.h:212-215):
//   ComputeNormalsOrganized            ps/algorithms.h:329-375
//   PrepareUnlabeledLabels             ps/segmentation.cc:21-26
//   FindSeedPointsFromPlaneSupport     ps/segmentation.h:190-230
//   SegmentRegions<PlanarRegion<>>     ps/segmentation.h:239-292 + ps/planar_region.h
//   SegmentRegions<ClusterRegion<>>    ps/segmentation.h:239-292 + ps/cluster_region.h
//   SlidingMeanShift                   ps/mean_shift_segmentation.h:207-330
// Clouds are duck-typed: any type with Data() / Rows() / Cols() whose elements are
// three packed floats (the reference's Cloud<eigenmath::Vector3f>, CloudView,
// CloudBuffer -- column-major, ps/cloud.h:37-41,74 -- or pcseg::CloudRef below)
// and Cloud<int> for labels.  Only PointT = NormalT = Vector3f is supported (the
// only instantiation the reference uses).  Config structs mirror the proto2
// messages of ps/region_segmentation_config.proto:22-72: unset optional fields
// take the defaults of ps/planar_region.h:93-121 / ps/cluster_region.h:53-63.
//   FindSeedPointsFromAverageNormals   ps/segmentation.h:137-184
//   FindSeedPointsFromLastPlanarRegions ps/planar_region.h:478-519
//   PlaneEstimator                     ps/plane_estimator.h:37-120
//   TransformInPlace (points/normals)  ps/multichannel_cloud.cc:53-68
//   DetectedObject geometry gathers    ps/detected_objects.cc:21-48, .h:62-74
// No CPU implementation lives here: every call goes to the CUDA library.
//
// Everything is declared in namespace mobility::b200, so this header can be included next to
// the reference's own planar_region.h / cluster_region.h / algorithms.h (whose class templates
// PlanarRegion<>, ClusterRegion<> and struct ComputeNormalsParams live in namespace mobility).
// A translation unit that does NOT include those headers can define
// PCSEG_B200_EXPORT_TO_MOBILITY before this one to have the same names visible as mobility::X.
// The result classes are not the reference's templates: centroid() / plane() return plain float
// arrays instead of eigenmath::Vector3f / Plane3f (Eigen is not a dependency of this library);
// ToReference() in INTEGRATION.md section 3 shows the conversion through the reference's own
// mutable_* accessors for callers of CreatePlanarDetectedObjectProto / ClassifyPlane.
#ifndef PCSEG_B200_HOST_H_
#define PCSEG_B200_HOST_H_

#include <cstdio>
#include <cstdlib>
#include <optional>
#include <utility>
#include <vector>

#include "pcseg_b200.h"

namespace mobility {
namespace b200 {

// CHECK-equivalent: the reference aborts on a failed precondition.
inline void CheckOk(pcseg_ctx* ctx, int status, const char* what) {
  if (status == PCSEG_OK) return;
  std::fprintf(stderr, "%s: CHECK failed: %s (%s)\n", what, pcseg_status_string(status),
               ctx ? pcseg_last_error(ctx) : "");
  std::abort();
}

// One context per GPU (RAII).  Not thread-safe per call, like the reference.
class Context {
 public:
  explicit Context(int device = 0) { CheckOk(nullptr, pcseg_create(device, &ctx_), "pcseg_create"); }
  ~Context() { pcseg_destroy(ctx_); }
  Context(const Context&) = delete;
  Context& operator=(const Context&) = delete;
  pcseg_ctx* get() const { return ctx_; }

 private:
  pcseg_ctx* ctx_ = nullptr;
};

// Minimal non-owning column-major view (index = col * rows + row) for callers
// that do not have the reference's Cloud types at hand.
template <typename T>
class CloudRef {
 public:
  CloudRef(T* data, int rows, int cols) : data_(data), rows_(rows), cols_(cols) {}
  T* Data() const { return data_; }
  int Rows() const { return rows_; }
  int Cols() const { return cols_; }
  int Size() const { return rows_ * cols_; }
  int LinearizeIndex(int row, int col) const { return col * rows_ + row; }

 private:
  T* data_;
  int rows_, cols_;
};

struct Float3 {
  float x, y, z;
};

template <typename CloudT>
inline const float* XyzData(const CloudT& c) {
  static_assert(sizeof(*c.Data()) == 3 * sizeof(float), "points/normals must be 3 packed floats");
  return reinterpret_cast<const float*>(c.Data());
}

struct RegionAccess;

// ps/algorithms.h:313-322
struct ComputeNormalsParams {
  float min_neighbor_distance = 0.1f;
  float max_neighbor_distance = 1.0f;
  bool include_diagonal_neighbors = true;
  int min_num_support_neighbors = 4;
};

// ps/region_segmentation_config.proto:42-72 (proto2 optional -> std::optional)
struct PlanarRegionConfig {
  std::optional<float> max_plane_distance, min_region_area;
  std::optional<int> min_region_inliers, plane_model_reestimation_period;
  std::optional<float> discontinuity_min_range, discontinuity_max_range, discontinuity_normal_angle_diff,
      discontinuity_z_diff, discontinuity_z_ratio;
};
// ps/region_segmentation_config.proto:22-39
struct ClusterRegionConfig {
  std::optional<int> min_region_inliers;
  std::optional<float> squared_distance_threshold;
  std::optional<int> half_search_window;
};

// Result of SegmentRegions<PlanarRegion<>>: accessors of ps/planar_region.h:427-450.
class PlanarRegion {
 public:
  using Config = PlanarRegionConfig;
  static pcseg_planar_config ApplyDefaultConfigValues(const Config& c) {  // ps/planar_region.h:93-121
    pcseg_planar_config d;
    pcseg_default_planar_config(&d);
    if (c.max_plane_distance) d.max_plane_distance = *c.max_plane_distance;
    if (c.min_region_area) d.min_region_area = *c.min_region_area;
    if (c.min_region_inliers) d.min_region_inliers = *c.min_region_inliers;
    if (c.plane_model_reestimation_period) d.plane_model_reestimation_period = *c.plane_model_reestimation_period;
    if (c.discontinuity_min_range) d.discontinuity_min_range = *c.discontinuity_min_range;
    if (c.discontinuity_max_range) d.discontinuity_max_range = *c.discontinuity_max_range;
    if (c.discontinuity_normal_angle_diff) d.discontinuity_normal_angle_diff = *c.discontinuity_normal_angle_diff;
    if (c.discontinuity_z_diff) d.discontinuity_z_diff = *c.discontinuity_z_diff;
    if (c.discontinuity_z_ratio) d.discontinuity_z_ratio = *c.discontinuity_z_ratio;
    return d;
  }
  // InputData of ps/planar_region.h:123-134; robot pose = quaternion (x, y, z, w) only (the
  // translation is never read on this path, :390).
  template <typename PointsT, typename NormalsT>
  struct InputData {
    float robot_q_xyzw[4];
    const PointsT& points;
    const NormalsT& normals;
  };
  const Float3& centroid() const { return centroid_; }
  const float* plane() const { return plane_; }  // normal xyz, offset (Plane3f::coeffs())
  const float* plane_estimator_state() const { return estimator_; }  // 6 second moments, sum xyz, weight
  int seed_point_index() const { return seed_point_index_; }
  float area() const { return area_; }
  const std::vector<int>& inlier_indices() const { return inlier_indices_; }
  const std::vector<int>& boundary_indices() const { return boundary_indices_; }
  const std::vector<int>& discontinuous_boundary_indices() const { return discontinuous_; }  // sorted set
  int label_id() const { return label_id_; }
  const std::vector<Float3>& projected_boundary_points() const { return hull_; }

 private:
  friend struct RegionAccess;
  int label_id_ = 0, seed_point_index_ = 0;
  Float3 centroid_{0, 0, 0};
  float plane_[4] = {1, 0, 0, 0}, area_ = 0.0f, estimator_[10] = {};
  std::vector<int> inlier_indices_, boundary_indices_, discontinuous_;
  std::vector<Float3> hull_;
};

// Result of SegmentRegions<ClusterRegion<>> / SlidingMeanShift: ps/cluster_region.h:161-168.
class ClusterRegion {
 public:
  using Config = ClusterRegionConfig;
  static pcseg_cluster_config ApplyDefaultConfigValues(const Config& c) {  // ps/cluster_region.h:53-63
    pcseg_cluster_config d;
    pcseg_default_cluster_config(&d);
    if (c.min_region_inliers) d.min_region_inliers = *c.min_region_inliers;
    if (c.squared_distance_threshold) d.squared_distance_threshold = *c.squared_distance_threshold;
    if (c.half_search_window) d.half_search_window = *c.half_search_window;
    return d;
  }
  const std::vector<int>& inlier_indices() const { return inlier_indices_; }
  int label_id() const { return label_id_; }
  const std::optional<Float3>& Seed() const { return seed_; }
  // ClusterUnorganized only (no counterpart in the reference): centroid and axis-aligned extent
  const std::optional<Float3>& Centroid() const { return centroid_; }
  const std::optional<Float3>& ExtentMin() const { return extent_min_; }
  const std::optional<Float3>& ExtentMax() const { return extent_max_; }

 private:
  friend struct RegionAccess;
  int label_id_ = 0;
  std::vector<int> inlier_indices_;
  std::optional<Float3> seed_, centroid_, extent_min_, extent_max_;
};

// Fills the region objects from a C-ABI result handle.
struct RegionAccess {
  static void Fill(const pcseg_regions* r, std::vector<PlanarRegion>* out) {
    const int n = pcseg_regions_count(r);
    const int* io = pcseg_regions_inlier_offsets(r);
    const int* ii = pcseg_regions_inlier_indices(r);
    const int* bo = pcseg_regions_boundary_offsets(r);
    const int* bi = pcseg_regions_boundary_indices(r);
    const int* dof = pcseg_regions_discontinuous_offsets(r);
    const int* di = pcseg_regions_discontinuous_indices(r);
    const int* ho = pcseg_regions_hull_offsets(r);
    const float* hp = pcseg_regions_hull_points(r);
    out->clear();  // regions->clear(), ps/segmentation.h:248
    out->resize(n);
    for (int k = 0; k < n; ++k) {
      PlanarRegion& p = (*out)[k];
      p.label_id_ = pcseg_regions_label_ids(r)[k];
      p.seed_point_index_ = pcseg_regions_seed_indices(r)[k];
      const float* c = pcseg_regions_centroids(r) + 3 * k;
      p.centroid_ = {c[0], c[1], c[2]};
      for (int j = 0; j < 4; ++j) p.plane_[j] = pcseg_regions_planes(r)[4 * k + j];
      for (int j = 0; j < 10; ++j) p.estimator_[j] = pcseg_regions_estimators(r)[10 * k + j];
      p.area_ = pcseg_regions_areas(r)[k];
      p.inlier_indices_.assign(ii + io[k], ii + io[k + 1]);
      p.boundary_indices_.assign(bi + bo[k], bi + bo[k + 1]);
      p.discontinuous_.assign(di + dof[k], di + dof[k + 1]);
      for (int j = ho[k]; j < ho[k + 1]; ++j) p.hull_.push_back({hp[3 * j], hp[3 * j + 1], hp[3 * j + 2]});
    }
  }
  static void Fill(const pcseg_regions* r, std::vector<ClusterRegion>* out, bool with_seed) {
    const int n = pcseg_regions_count(r);
    const int* io = pcseg_regions_inlier_offsets(r);
    const int* ii = pcseg_regions_inlier_indices(r);
    out->clear();
    out->resize(n);
    for (int k = 0; k < n; ++k) {
      ClusterRegion& c = (*out)[k];
      c.label_id_ = pcseg_regions_label_ids(r)[k];
      c.inlier_indices_.assign(ii + io[k], ii + io[k + 1]);
      if (with_seed) {
        const float* s = pcseg_regions_seed_points(r) + 3 * k;
        c.seed_ = Float3{s[0], s[1], s[2]};
      }
    }
  }
  static void FillStats(const pcseg_regions* r, std::vector<ClusterRegion>* out) {
    const float* cen = pcseg_regions_centroids(r);
    const float* box = pcseg_regions_bboxes(r);
    for (size_t k = 0; k < out->size(); ++k) {
      ClusterRegion& c = (*out)[k];
      c.centroid_ = Float3{cen[3 * k], cen[3 * k + 1], cen[3 * k + 2]};
      c.extent_min_ = Float3{box[6 * k], box[6 * k + 1], box[6 * k + 2]};
      c.extent_max_ = Float3{box[6 * k + 3], box[6 * k + 4], box[6 * k + 5]};
    }
  }
};

// ps/algorithms.h:329-375 (both overloads).  Only the translation of
// point_cloud_pose_sensor is read by the reference (:355).
template <typename PointsT, typename NormalsT>
void ComputeNormalsOrganized(Context& ctx, const double sensor_translation[3], const PointsT& points,
                             NormalsT* normals, const std::pair<int, int>& row_range,
                             const std::pair<int, int>& col_range,
                             const ComputeNormalsParams& params = ComputeNormalsParams()) {
  pcseg_normals_params p{params.min_neighbor_distance, params.max_neighbor_distance,
                         params.include_diagonal_neighbors ? 1 : 0, params.min_num_support_neighbors};
  CheckOk(ctx.get(),
                pcseg_normals_organized(ctx.get(), XyzData(points), points.Rows(), points.Cols(),
                                        sensor_translation, &p, row_range.first, row_range.second,
                                        col_range.first, col_range.second,
                                        const_cast<float*>(XyzData(*normals)), nullptr),
                "ComputeNormalsOrganized");
}
template <typename PointsT, typename NormalsT>
void ComputeNormalsOrganized(Context& ctx, const double sensor_translation[3], const PointsT& points,
                             NormalsT* normals, const ComputeNormalsParams& params = ComputeNormalsParams()) {
  ComputeNormalsOrganized(ctx, sensor_translation, points, normals, {0, points.Rows()}, {0, points.Cols()}, params);
}

// ps/segmentation.cc:21-26
template <typename PointsT>
std::vector<int> PrepareUnlabeledLabels(Context& ctx, const PointsT& point_cloud) {
  std::vector<int> labels(static_cast<size_t>(point_cloud.Rows()) * point_cloud.Cols());
  CheckOk(ctx.get(), pcseg_prepare_unlabeled_labels(ctx.get(), labels.data(), (int)labels.size()),
                "PrepareUnlabeledLabels");
  return labels;
}

// ps/segmentation.h:190-230
template <typename PointsT, typename NormalsT>
std::vector<int> FindSeedPointsFromPlaneSupport(Context& ctx, const PointsT& points, const NormalsT& normals,
                                                int neighborhood_size = 9, float max_plane_distance = 0.05f,
                                                int min_num_support_points = 12) {
  if (points.Cols() != normals.Cols() || points.Rows() != normals.Rows()) CheckOk(ctx.get(), PCSEG_ERR_DIMS, "FindSeedPointsFromPlaneSupport");
  std::vector<int> seeds(static_cast<size_t>(points.Rows()) * points.Cols());
  int n = 0;
  CheckOk(ctx.get(),
                pcseg_seeds_plane_support(ctx.get(), XyzData(points), XyzData(normals), points.Rows(),
                                          points.Cols(), neighborhood_size, max_plane_distance,
                                          min_num_support_points, seeds.data(), (int)seeds.size(), &n),
                "FindSeedPointsFromPlaneSupport");
  seeds.resize(n);
  return seeds;
}

// ps/segmentation.h:239-292 with RegionT = PlanarRegion<>.  labels: Cloud<int>-like, in/out.
template <typename PointsT, typename NormalsT, typename LabelsT>
void SegmentRegions(Context& ctx, const PlanarRegion::InputData<PointsT, NormalsT>& input,
                    const std::vector<int>& seed_point_indices, const PlanarRegionConfig& region_config,
                    std::vector<PlanarRegion>* regions, LabelsT* labels, int initial_id_offset = 0) {
  if (!regions || !labels) CheckOk(ctx.get(), PCSEG_ERR_NULL, "SegmentRegions<PlanarRegion>");  // CHECK(regions)
  // CheckInputDataAndLabels, ps/planar_region.h:257-268
  if (input.points.Cols() != input.normals.Cols() || input.points.Rows() != input.normals.Rows() ||
      input.points.Cols() != labels->Cols() || input.points.Rows() != labels->Rows())
    CheckOk(ctx.get(), PCSEG_ERR_DIMS, "SegmentRegions<PlanarRegion>");
  const pcseg_planar_config cfg = PlanarRegion::ApplyDefaultConfigValues(region_config);
  pcseg_regions* r = nullptr;
  CheckOk(ctx.get(),
                pcseg_segment_planar(ctx.get(), XyzData(input.points), XyzData(input.normals),
                                     input.points.Rows(), input.points.Cols(), input.robot_q_xyzw,
                                     seed_point_indices.data(), (int)seed_point_indices.size(), &cfg,
                                     initial_id_offset, labels->Data(), &r),
                "SegmentRegions<PlanarRegion>");
  RegionAccess::Fill(r, regions);
  pcseg_regions_free(r);
}

// ps/segmentation.h:239-292 with RegionT = ClusterRegion<> (InputData = the point cloud).
template <typename PointsT, typename LabelsT>
void SegmentRegions(Context& ctx, const PointsT& points, const std::vector<int>& seed_point_indices,
                    const ClusterRegionConfig& region_config, std::vector<ClusterRegion>* regions, LabelsT* labels,
                    int initial_id_offset = 0) {
  if (!regions || !labels) CheckOk(ctx.get(), PCSEG_ERR_NULL, "SegmentRegions<ClusterRegion>");
  if (points.Cols() != labels->Cols() || points.Rows() != labels->Rows())  // ps/cluster_region.h:152-159
    CheckOk(ctx.get(), PCSEG_ERR_DIMS, "SegmentRegions<ClusterRegion>");
  const pcseg_cluster_config cfg = ClusterRegion::ApplyDefaultConfigValues(region_config);
  pcseg_regions* r = nullptr;
  CheckOk(ctx.get(),
                pcseg_segment_cluster(ctx.get(), XyzData(points), points.Rows(), points.Cols(),
                                      seed_point_indices.data(), (int)seed_point_indices.size(), &cfg,
                                      initial_id_offset, labels->Data(), &r),
                "SegmentRegions<ClusterRegion>");
  RegionAccess::Fill(r, regions, false);
  pcseg_regions_free(r);
}

// No counterpart in the reference (it clusters organized grids only): fixed-radius Euclidean clustering of an
// UNORGANIZED cloud of n points (xyz: 3 floats each) through a voxel hash, pcseg_cluster_unorganized.
// labels (n ints, in/out): points with another value than kUnlabeled (-1) do not take part.
inline void ClusterUnorganized(Context& ctx, const float* xyz, int n, float radius, int min_region_inliers,
                               int initial_id_offset, int* labels, std::vector<ClusterRegion>* regions) {
  if (!regions || !labels || !xyz) CheckOk(ctx.get(), PCSEG_ERR_NULL, "ClusterUnorganized");
  pcseg_voxel_cluster_config cfg;
  cfg.radius = radius;
  cfg.min_region_inliers = min_region_inliers;
  pcseg_regions* r = nullptr;
  CheckOk(ctx.get(), pcseg_cluster_unorganized(ctx.get(), xyz, n, &cfg, initial_id_offset, labels, &r), "ClusterUnorganized");
  RegionAccess::Fill(r, regions, false);
  RegionAccess::FillStats(r, regions);
  pcseg_regions_free(r);
}

// ps/mean_shift_segmentation.h:207-211.  The sort of :276-280 is std::sort of this toolchain's
// standard library, as in the reference (quirk Q7).
template <typename PointsT, typename LabelsT>
void SlidingMeanShift(Context& ctx, const PointsT& point_cloud, const ClusterRegionConfig& config,
                      int iterations, int initial_region_id_offset, LabelsT* labels,
                      std::vector<ClusterRegion>* regions) {
  if (!regions || !labels) CheckOk(ctx.get(), PCSEG_ERR_NULL, "SlidingMeanShift");  // CHECK_NE(.., nullptr)
  if (point_cloud.Cols() != labels->Cols() || point_cloud.Rows() != labels->Rows())
    CheckOk(ctx.get(), PCSEG_ERR_DIMS, "SlidingMeanShift");
  const pcseg_cluster_config cfg = ClusterRegion::ApplyDefaultConfigValues(config);
  pcseg_regions* r = nullptr;
  CheckOk(ctx.get(),
                pcseg_mean_shift(ctx.get(), XyzData(point_cloud), point_cloud.Rows(), point_cloud.Cols(), &cfg,
                                 iterations, initial_region_id_offset, 0, labels->Data(), &r),
                "SlidingMeanShift");
  RegionAccess::Fill(r, regions, true);
  pcseg_regions_free(r);
}


// ps/segmentation.h:137-184 (quirk Q1 literal: rows <= cols required, see the C ABI)
template <typename PointsT, typename NormalsT>
std::vector<int> FindSeedPointsFromAverageNormals(Context& ctx, const PointsT& points, const NormalsT& normals,
                                                  int neighborhood_size = 5, int min_num_valid_normals = 16,
                                                  float min_avg_normal_length = 0.9f) {
  if (points.Cols() != normals.Cols() || points.Rows() != normals.Rows()) CheckOk(ctx.get(), PCSEG_ERR_DIMS, "FindSeedPointsFromAverageNormals");
  std::vector<int> seeds(static_cast<size_t>(points.Rows()) * points.Cols());
  int n = 0;
  CheckOk(ctx.get(),
          pcseg_seeds_average_normals(ctx.get(), XyzData(points), XyzData(normals), points.Rows(), points.Cols(),
                                      neighborhood_size, min_num_valid_normals, min_avg_normal_length, seeds.data(),
                                      (int)seeds.size(), &n),
          "FindSeedPointsFromAverageNormals");
  seeds.resize(n);
  return seeds;
}

// ps/planar_region.h:478-519.  current_pose_prev: quaternion (x, y, z, w) + translation, an
// eigenmath::Pose3f.  Found seeds are appended to *seed_point_indices, as in the reference.
template <typename PointsT, typename NormalsT>
void FindSeedPointsFromLastPlanarRegions(Context& ctx, const PointsT& points, const NormalsT& normals,
                                         const std::vector<PlanarRegion>& prev_regions, const float q_xyzw[4],
                                         const float translation[3], float max_distance,
                                         float max_normal_difference_angle, std::vector<int>* seed_point_indices) {
  if (!seed_point_indices) CheckOk(ctx.get(), PCSEG_ERR_NULL, "FindSeedPointsFromLastPlanarRegions");
  const int n_prev = (int)prev_regions.size();
  std::vector<int> counts(n_prev);
  std::vector<float> cent(3 * (size_t)n_prev), nrm(3 * (size_t)n_prev);
  for (int i = 0; i < n_prev; ++i) {
    counts[i] = (int)prev_regions[i].inlier_indices().size();
    const Float3& c = prev_regions[i].centroid();
    cent[3 * i] = c.x; cent[3 * i + 1] = c.y; cent[3 * i + 2] = c.z;
    for (int j = 0; j < 3; ++j) nrm[3 * i + j] = prev_regions[i].plane()[j];
  }
  int n = (int)seed_point_indices->size();
  seed_point_indices->resize((size_t)n + n_prev);
  CheckOk(ctx.get(),
          pcseg_seeds_last_planar_regions(ctx.get(), XyzData(points), XyzData(normals), points.Rows() * points.Cols(),
                                          n_prev, counts.data(), cent.data(), nrm.data(), q_xyzw, translation,
                                          max_distance, max_normal_difference_angle, seed_point_indices->data(),
                                          (int)seed_point_indices->size(), &n),
          "FindSeedPointsFromLastPlanarRegions");
  seed_point_indices->resize(n);
}

// ps/plane_estimator.h:37-120.  The estimator keeps the operations since the last Clear() and
// replays them on the device when a result is asked for (lazy recomputation and the kept normal
// orientation are the reference's, ps/plane_estimator.cc:85-126,184-229).  Bulk fits of many index
// sets go through FitPlanes below (one launch for all regions).
class PlaneEstimator {
 public:
  explicit PlaneEstimator(Context& ctx) : ctx_(&ctx) { Clear(); }
  void Clear() { ops_.assign(1, 4); args_.assign(4, 0.0f); dirty_ = true; }
  void AddPoint(const Float3& p) { Push(0, p.x, p.y, p.z, 0.0f); }
  void AddPoint(const Float3& p, float weight) { Push(1, p.x, p.y, p.z, weight); }
  template <typename PointsT>
  void AddPoints(const PointsT& points, const std::vector<int>& indices) {  // ps/plane_estimator.h:97-103
    const float* d = XyzData(points);
    for (int i : indices) Push(0, d[3 * (size_t)i], d[3 * (size_t)i + 1], d[3 * (size_t)i + 2], 0.0f);
  }
  void SetNormalOrientation(const Float3& n) { Push(2, n.x, n.y, n.z, 0.0f); }
  bool PlaneValid() { return Query()[0] != 0.0f; }
  float Curvature() { return Query()[1]; }
  Float3 Centroid() { const float* q = Query(); return Float3{q[2], q[3], q[4]}; }
  const float* Plane() { return Query() + 5; }  // normal xyz, offset
  float CumulativeWeight() { return Query()[9]; }

 private:
  void Push(int op, float a, float b, float c, float d) {
    ops_.push_back(op);
    args_.insert(args_.end(), {a, b, c, d});
    dirty_ = true;
  }
  const float* Query() {
    if (dirty_) {
      std::vector<int> ops(ops_);
      std::vector<float> args(args_);
      ops.push_back(3);
      args.insert(args.end(), {0.0f, 0.0f, 0.0f, 0.0f});
      int nq = 0;
      CheckOk(ctx_->get(),
              pcseg_plane_estimator_replay(ctx_->get(), ops.data(), args.data(), (int)ops.size(), nullptr, 0, row_, 1, &nq),
              "PlaneEstimator");
      dirty_ = false;
    }
    return row_;
  }
  Context* ctx_;
  std::vector<int> ops_;
  std::vector<float> args_;
  float row_[10] = {};
  bool dirty_ = true;
};

// Least-squares plane fits of many index sets in one launch (CSR: offsets[n_regions + 1]): per
// region Clear(); AddPoints(points, indices) in list order; Plane() / Centroid() / Curvature().
// query_out: 10 floats per region {valid, curvature, centroid xyz, normal xyz, offset, weight}.
template <typename PointsT>
void FitPlanes(Context& ctx, const PointsT& points, const std::vector<int>& offsets, const std::vector<int>& indices,
               std::vector<float>* query_out, std::vector<float>* state_out = nullptr) {
  const int n = (int)offsets.size() - 1;
  query_out->assign(10 * (size_t)(n > 0 ? n : 0), 0.0f);
  if (state_out) state_out->assign(10 * (size_t)(n > 0 ? n : 0), 0.0f);
  if (n <= 0) return;
  CheckOk(ctx.get(),
          pcseg_fit_planes(ctx.get(), XyzData(points), points.Rows() * points.Cols(), offsets.data(), indices.data(), n,
                           nullptr, nullptr, query_out->data(), state_out ? state_out->data() : nullptr),
          "FitPlanes");
}

// MultichannelCloud::TransformInPlace for the Points / Normals channels, ps/multichannel_cloud.cc:53-68.
template <typename PointsT, typename NormalsT>
void TransformInPlace(Context& ctx, const double q_xyzw[4], const double translation[3], PointsT* points,
                      NormalsT* normals) {
  const int n = points ? points->Rows() * points->Cols() : (normals ? normals->Rows() * normals->Cols() : 0);
  CheckOk(ctx.get(),
          pcseg_transform_in_place(ctx.get(), points ? const_cast<float*>(XyzData(*points)) : nullptr,
                                   normals ? const_cast<float*>(XyzData(*normals)) : nullptr, n, q_xyzw, translation),
          "TransformInPlace");
}

// The geometry CreatePlanarDetectedObjectProto / CreateClusterDetectedObjectProto put into the proto
// (ps/detected_objects.cc:21-48, ps/detected_objects.h:62-74): the inlier points in inlier order and,
// for planes, the positions in that list of the discontinuous boundary points.  Serialisation
// stays with the caller.
struct DetectedObjectGeometry {
  std::vector<Float3> points;
  std::vector<int> discontinuous_positions;
};
template <typename PointsT>
DetectedObjectGeometry CreatePlanarDetectedObjectGeometry(Context& ctx, const PointsT& points, const PlanarRegion& region) {
  DetectedObjectGeometry g;
  const std::vector<int>& idx = region.inlier_indices();
  const std::vector<int>& disc = region.discontinuous_boundary_indices();
  g.points.resize(idx.size());
  g.discontinuous_positions.resize(disc.size());
  int np = 0;
  CheckOk(ctx.get(),
          pcseg_gather_points(ctx.get(), XyzData(points), points.Rows() * points.Cols(), idx.data(), (int)idx.size(),
                              reinterpret_cast<float*>(g.points.data()), disc.data(), (int)disc.size(),
                              g.discontinuous_positions.data(), &np),
          "CreatePlanarDetectedObjectGeometry");
  g.discontinuous_positions.resize(np);
  return g;
}
template <typename PointsT>
DetectedObjectGeometry CreateClusterDetectedObjectGeometry(Context& ctx, const PointsT& points, const ClusterRegion& region) {
  DetectedObjectGeometry g;
  const std::vector<int>& idx = region.inlier_indices();
  g.points.resize(idx.size());
  int np = 0;
  CheckOk(ctx.get(),
          pcseg_gather_points(ctx.get(), XyzData(points), points.Rows() * points.Cols(), idx.data(), (int)idx.size(),
                              reinterpret_cast<float*>(g.points.data()), nullptr, 0, nullptr, &np),
          "CreateClusterDetectedObjectGeometry");
  return g;
}

}  // namespace b200
}  // namespace mobility

#ifdef PCSEG_B200_EXPORT_TO_MOBILITY
namespace mobility {
using namespace b200;  // NOLINT: opt-in for translation units without the reference headers
}
#endif
#endif  // PCSEG_B200_HOST_H_
