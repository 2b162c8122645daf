// TEST INFRASTRUCTURE ONLY.  Hand-written stand-in for the protoc output of
// ps/plane_estimator.proto:22-32 (proto3: three repeated floats and one float).
#ifndef ORACLE_SHIMS_PS_PLANE_ESTIMATOR_PB_H_
#define ORACLE_SHIMS_PS_PLANE_ESTIMATOR_PB_H_
#include <vector>
namespace mobility {
class PlaneEstimatorProto {
 public:
#define SHIM_PROTO_REPEATED(name)                         \
  float name(int i) const { return name##_[i]; }          \
  int name##_size() const { return static_cast<int>(name##_.size()); } \
  void add_##name(float v) { name##_.push_back(v); }      \
  void clear_##name() { name##_.clear(); }
  SHIM_PROTO_REPEATED(cumulative_centroid)
  SHIM_PROTO_REPEATED(covariance_accumulator)
  SHIM_PROTO_REPEATED(normal)
#undef SHIM_PROTO_REPEATED
  float cumulative_weights() const { return cumulative_weights_; }
  void set_cumulative_weights(float v) { cumulative_weights_ = v; }
 private:
  std::vector<float> cumulative_centroid_, covariance_accumulator_, normal_;
  float cumulative_weights_ = 0.0f;
};
}  // namespace mobility
#endif
