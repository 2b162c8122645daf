// Compile test: the host-side mirror (namespace mobility::b200) included in ONE translation
// unit together with the reference's own region headers (namespace mobility), as a caller
// migrating call site by call site would have it.  Built only where /root/reference exists,
// against the oracle's shim headers for Eigen / eigenmath / abseil / protobuf.
#include "pointcloud_segmentation/algorithms.h"
#include "pointcloud_segmentation/cluster_region.h"
#include "pointcloud_segmentation/planar_region.h"
#include "pointcloud_segmentation/segmentation.h"

#include <type_traits>

#include "pcseg_host.h"

// the reference's types and the mirror's types are distinct and both usable
static_assert(!std::is_same<mobility::PlanarRegion<>, mobility::b200::PlanarRegion>::value, "");
static_assert(!std::is_same<mobility::ComputeNormalsParams, mobility::b200::ComputeNormalsParams>::value, "");

// a b200 result converted into the reference's own region type
// through its mutable accessors, for CreatePlanarDetectedObjectProto / ClassifyPlane callers.
mobility::PlanarRegion<> ToReference(const mobility::b200::PlanarRegion& r, const mobility::PlanarRegionConfigProto& config) {
  mobility::PlanarRegion<> out(config);
  *out.MutableInlierIndicesForTestOnly() = r.inlier_indices();  // the only mutable access the reference offers
  const mobility::b200::Float3& c = r.centroid();
  *out.mutable_centroid() = eigenmath::Vector3f(c.x, c.y, c.z);
  const float* p = r.plane();
  *out.mutable_plane() = eigenmath::Plane3f(eigenmath::Vector3f(p[0], p[1], p[2]), p[3]);
  *out.mutable_seed_point_index() = r.seed_point_index();
  *out.mutable_label_id() = r.label_id();
  return out;
}

int main() { return 0; }
