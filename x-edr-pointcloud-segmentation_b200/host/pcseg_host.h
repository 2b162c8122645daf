// pcseg_host.h -- C++17 host-side mirror of the reference's segmentation API for
// the hot path, on top of the C ABI (include/pcseg_b200.h).
//
// Same names, argument meaning and error behaviour as the reference
// (namespace mobility; precondition failure = process abort like absl CHECK,
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
// No CPU implementation lives here: every call goes to the CUDA library.
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

}  // namespace b200

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
  const b200::Float3& centroid() const { return centroid_; }
  const float* plane() const { return plane_; }  // normal xyz, offset (Plane3f::coeffs())
  const float* plane_estimator_state() const { return estimator_; }  // 6 second moments, sum xyz, weight
  int seed_point_index() const { return seed_point_index_; }
  float area() const { return area_; }
  const std::vector<int>& inlier_indices() const { return inlier_indices_; }
  const std::vector<int>& boundary_indices() const { return boundary_indices_; }
  const std::vector<int>& discontinuous_boundary_indices() const { return discontinuous_; }  // sorted set
  int label_id() const { return label_id_; }
  const std::vector<b200::Float3>& projected_boundary_points() const { return hull_; }

 private:
  template <typename R, typename I, typename C>
  friend void SegmentRegions(b200::Context&, const I&, const std::vector<int>&, const C&, std::vector<R>*,
                             b200::CloudRef<int>, int);
  friend struct RegionAccess;
  int label_id_ = 0, seed_point_index_ = 0;
  b200::Float3 centroid_{0, 0, 0};
  float plane_[4] = {1, 0, 0, 0}, area_ = 0.0f, estimator_[10] = {};
  std::vector<int> inlier_indices_, boundary_indices_, discontinuous_;
  std::vector<b200::Float3> hull_;
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
  const std::optional<b200::Float3>& Seed() const { return seed_; }

 private:
  friend struct RegionAccess;
  int label_id_ = 0;
  std::vector<int> inlier_indices_;
  std::optional<b200::Float3> seed_;
};
