// TEST INFRASTRUCTURE ONLY.  Hand-written stand-in for the protoc output of
// ps/region_segmentation_config.proto:22-72 (protoc / libprotobuf are not in this image):
// proto2 optional fields with has_x() / x() / set_x() / clear_x().
#ifndef ORACLE_SHIMS_PS_REGION_SEGMENTATION_CONFIG_PB_H_
#define ORACLE_SHIMS_PS_REGION_SEGMENTATION_CONFIG_PB_H_
#include <cstdint>
namespace mobility {
#define SHIM_PROTO_OPTIONAL(type, name)           \
 public:                                          \
  bool has_##name() const { return has_##name##_; } \
  type name() const { return name##_; }           \
  void set_##name(type v) {                       \
    name##_ = v;                                  \
    has_##name##_ = true;                         \
  }                                               \
  void clear_##name() {                           \
    name##_ = type();                             \
    has_##name##_ = false;                        \
  }                                               \
 private:                                         \
  type name##_ = type();                          \
  bool has_##name##_ = false;

class ClusterRegionConfigProto {
 public:
  enum ClusterMethod { NEAREST_NEIGHBOR_DEFAULT = 0, MEAN_SHIFT = 1 };
  SHIM_PROTO_OPTIONAL(int32_t, min_region_inliers)
  SHIM_PROTO_OPTIONAL(float, squared_distance_threshold)
  SHIM_PROTO_OPTIONAL(int32_t, half_search_window)
  SHIM_PROTO_OPTIONAL(int, cluster_method)
};

class PlanarRegionConfigProto {
  SHIM_PROTO_OPTIONAL(float, max_distance_for_seed_point)
  SHIM_PROTO_OPTIONAL(float, max_normal_difference_angle_for_seed_point)
  SHIM_PROTO_OPTIONAL(float, max_plane_distance)
  SHIM_PROTO_OPTIONAL(float, min_region_area)
  SHIM_PROTO_OPTIONAL(int32_t, min_region_inliers)
  SHIM_PROTO_OPTIONAL(int32_t, plane_model_reestimation_period)
  SHIM_PROTO_OPTIONAL(float, discontinuity_min_range)
  SHIM_PROTO_OPTIONAL(float, discontinuity_max_range)
  SHIM_PROTO_OPTIONAL(float, discontinuity_normal_angle_diff)
  SHIM_PROTO_OPTIONAL(float, discontinuity_z_diff)
  SHIM_PROTO_OPTIONAL(float, discontinuity_z_ratio)
};
#undef SHIM_PROTO_OPTIONAL
}  // namespace mobility
#endif
