// ref_driver.cc -- TEST INFRASTRUCTURE ONLY (never shipped, never linked into
// libpcseg_b200.so).  Exposes the REFERENCE's own hot-path code -- compiled in
// place from $(REF)/pointcloud_segmentation/*.{h,cc} against the shim headers
// in oracle/shims/ (see oracle/Makefile) -- through the same C interface as the
// restatement (oracle/pcseg_oracle.h), so that tests/test_oracle_vs_ref.py can
// run both on the same inputs.  Everything below is glue: the algorithms are
// the reference's (ps/segmentation.h:239-292, ps/planar_region.h,
// ps/cluster_region.h, ps/plane_estimator.cc, ps/algorithms.h,
// ps/mean_shift_segmentation.h).  The only deviation is patch P1:
// the Makefile generates _ref/patched/pointcloud_segmentation/algorithms.h with
// the convex-hull output buffer sized before it is written (ps/algorithms.h:543).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <utility>
#include <vector>

#include "pcseg_oracle.h"
#include "pointcloud_segmentation/algorithms.h"
#include "pointcloud_segmentation/cluster_region.h"
#include "pointcloud_segmentation/mean_shift_segmentation.h"
#include "pointcloud_segmentation/planar_region.h"
#include "pointcloud_segmentation/plane_estimator.h"
#include "pointcloud_segmentation/segmentation.h"

using eigenmath::Vector3f;
using mobility::Cloud;
using mobility::CloudBuffer;
using mobility::CloudView;

static_assert(sizeof(Vector3f) == 12, "Vector3f must alias float[3]");

struct oracle_regions {
  std::vector<int> inlier_offsets{0};
  std::vector<int> inlier_indices;
  std::vector<int> label_ids;
  std::vector<int> seed_indices;
  std::vector<float> planes;
  std::vector<float> centroids;
  std::vector<float> areas;
  std::vector<float> estimators;
  std::vector<int> boundary_offsets{0};
  std::vector<int> boundary_indices;
  std::vector<int> disc_offsets{0};
  std::vector<int> disc_indices;
  std::vector<int> hull_offsets{0};
  std::vector<float> hull_points;
  std::vector<float> seed_points;
};

namespace {

CloudView<Vector3f> PointsView(const float* xyz, int rows, int cols) {
  return CloudView<Vector3f>(reinterpret_cast<Vector3f*>(const_cast<float*>(xyz)), rows, cols);
}
CloudView<const Vector3f> ConstView(const float* xyz, int rows, int cols) {
  return CloudView<const Vector3f>(reinterpret_cast<const Vector3f*>(xyz), rows, cols);
}

void AppendInliers(oracle_regions* out, const std::vector<int>& inl, int label) {
  out->inlier_indices.insert(out->inlier_indices.end(), inl.begin(), inl.end());
  out->inlier_offsets.push_back(static_cast<int>(out->inlier_indices.size()));
  out->label_ids.push_back(label);
}

template <typename F>
int Guarded(F&& f) {
  try {
    f();
    return 0;
  } catch (const shim::CheckFailure&) {
    return 1;  // the reference would have aborted on a CHECK
  }
}

}  // namespace

extern "C" {

void oracle_set_native_libm(int on) { shim::native_libm() = on; }

int oracle_normals_organized(const float* xyz, int rows, int cols, const double sensor_t[3],
                             const oracle_normals_params* p, int row_begin, int row_end, int col_begin,
                             int col_end, float* normals, float* curvature_or_null) {
  if (curvature_or_null) return 2;  // the reference has no curvature output (extension of the port)
  return Guarded([&] {
    auto points = PointsView(xyz, rows, cols);
    CloudView<Vector3f> nrm(reinterpret_cast<Vector3f*>(normals), rows, cols);
    eigenmath::Pose3d pose(eigenmath::Quaterniond::Identity(),
                           eigenmath::Vector3d(sensor_t[0], sensor_t[1], sensor_t[2]));
    mobility::ComputeNormalsParams params;
    params.min_neighbor_distance = p->min_neighbor_distance;
    params.max_neighbor_distance = p->max_neighbor_distance;
    params.include_diagonal_neighbors = p->include_diagonal_neighbors != 0;
    params.min_num_support_neighbors = p->min_num_support_neighbors;
    mobility::ComputeNormalsOrganized(pose, static_cast<const Cloud<Vector3f>&>(points),
                                      static_cast<Cloud<Vector3f>*>(&nrm), {row_begin, row_end},
                                      {col_begin, col_end}, params);
  });
}

int oracle_find_normal_support_neighbors(const float* xyz, int rows, int cols, int col, int row, float min_d,
                                         float max_d, int diag, int* out) {
  auto points = PointsView(xyz, rows, cols);
  std::vector<int> idx;
  const int n = mobility::detail::FindNormalSupportNeighbors(static_cast<const Cloud<Vector3f>&>(points), col,
                                                             row, min_d, max_d, diag != 0, &idx);
  for (size_t i = 0; i < idx.size(); ++i) out[i] = idx[i];
  return n;
}

int oracle_plane_estimator_replay(const int* ops, const float* args, int n_ops, float* out, int out_cap) {
  return oracle_plane_estimator_replay2(ops, args, n_ops, nullptr, 0, out, out_cap);
}

int oracle_plane_estimator_replay2(const int* ops, const float* args, int n_ops, const double* poses, int n_poses,
                                   float* out, int out_cap) {
  mobility::PlaneEstimator slots[4];
  int cur = 0;
  int nq = 0;
  auto pose = [&](int k) {
    const double* q = poses + 7 * k;
    return eigenmath::Pose3d(eigenmath::Quaterniond(q[3], q[0], q[1], q[2]), eigenmath::Vector3d(q[4], q[5], q[6]));
  };
  auto weight = [](const mobility::PlaneEstimator& e) {
    mobility::PlaneEstimatorProto proto;
    e.ToProto(&proto);
    return proto.cumulative_weights();
  };
  for (int i = 0; i < n_ops; ++i) {
    mobility::PlaneEstimator& est = slots[cur];
    const float* a = args + 4 * i;
    const Vector3f p(a[0], a[1], a[2]);
    switch (ops[i]) {
      case 5: {
        const int k = static_cast<int>(a[0]);
        if (k < 0 || k >= 4) return -2;
        cur = k;
        break;
      }
      case 6: {
        const int k = static_cast<int>(a[0]);
        if (k < 0 || k >= 4 || k == cur) return -2;
        est.Merge(slots[k]);
        break;
      }
      case 7: {
        const int pi = static_cast<int>(a[0]);
        if (pi < 0 || pi >= n_poses) return -2;
        if (!(weight(est) > 0.0f)) return -3;  // the reference CHECK-aborts
        est.Transform(pose(pi));
        break;
      }
      case 8: {
        const int k = static_cast<int>(a[0]), pi = static_cast<int>(a[1]);
        if (k < 0 || k >= 4 || k == cur || pi < 0 || pi >= n_poses) return -2;
        if (!(weight(slots[k]) > 0.0f)) return -3;
        est.Merge(slots[k], pose(pi));
        break;
      }
      case 0: est.AddPoint(p); break;
      case 1: est.AddPoint(p, a[3]); break;
      case 2: est.SetNormalOrientation(p); break;
      case 4: est.Clear(); break;
      case 3: {
        if (nq >= out_cap) return nq;
        float* o = out + 10 * nq++;
        mobility::PlaneEstimatorProto proto;
        est.ToProto(&proto);
        o[0] = est.PlaneValid() ? 1.0f : 0.0f;
        o[1] = est.Curvature();
        const Vector3f c = est.Centroid();
        o[2] = c.x(); o[3] = c.y(); o[4] = c.z();
        const auto& pl = est.Plane();
        o[5] = pl.normal().x(); o[6] = pl.normal().y(); o[7] = pl.normal().z();
        o[8] = pl.offset();
        o[9] = proto.cumulative_weights();
        break;
      }
      default: return -1;
    }
  }
  return nq;
}

void oracle_eig3(const float* a, float* evals, float* evecs) {
  eigenmath::Matrix3f m;
  m(0, 0) = a[0]; m(1, 0) = a[1]; m(2, 0) = a[2]; m(1, 1) = a[3]; m(2, 1) = a[4]; m(2, 2) = a[5];
  m(0, 1) = a[1]; m(0, 2) = a[2]; m(1, 2) = a[4];
  Eigen::SelfAdjointEigenSolver<eigenmath::Matrix3f> s;
  s.computeDirect(m);
  for (int i = 0; i < 3; ++i) evals[i] = s.eigenvalues()(i);
  for (int i = 0; i < 9; ++i) evecs[i] = s.eigenvectors().coeff(i);
}

float oracle_polygon_area(const float* xyz, int n_points, const int* indices, int n_indices,
                          float* cumulative_normal_or_null) {
  CloudView<Vector3f> points(reinterpret_cast<Vector3f*>(const_cast<float*>(xyz)), 1, n_points);
  const Cloud<Vector3f>& pc = points;
  Eigen::Vector3f n;
  float area;
  if (indices) {
    const mobility::Indices idx(indices, indices + n_indices);
    n = mobility::CumulativePolygonNormal(pc, idx);
    area = mobility::PolygonArea(pc, idx);
  } else {
    n = mobility::CumulativePolygonNormal(pc, mobility::NoIndices());
    area = mobility::PolygonArea(pc, mobility::NoIndices());
  }
  if (cumulative_normal_or_null) {
    cumulative_normal_or_null[0] = n.x();
    cumulative_normal_or_null[1] = n.y();
    cumulative_normal_or_null[2] = n.z();
  }
  return area;
}

void oracle_prepare_unlabeled_labels(int* labels, int n) {
  // ps/segmentation.cc:21-26 on a 1 x n cloud (only Rows()/Cols() of the argument are read)
  CloudView<Vector3f> pc(nullptr, 1, n);
  CloudBuffer<int> l = mobility::PrepareUnlabeledLabels(pc);
  std::memcpy(labels, l.Data(), sizeof(int) * static_cast<size_t>(n));
}

int oracle_segment_planar(const float* xyz, const float* normals, int rows, int cols,
                          const float robot_q_xyzw[4], const int* seeds, int n_seeds,
                          const oracle_planar_config* cfg, int id_offset, int* labels, oracle_regions** out) {
  auto* res = new oracle_regions;
  *out = res;
  return Guarded([&] {
    using Region = mobility::PlanarRegion<>;
    mobility::PlanarRegionConfigProto config;
    config.set_max_plane_distance(cfg->max_plane_distance);
    config.set_min_region_area(cfg->min_region_area);
    config.set_min_region_inliers(cfg->min_region_inliers);
    config.set_plane_model_reestimation_period(cfg->plane_model_reestimation_period);
    config.set_discontinuity_min_range(cfg->discontinuity_min_range);
    config.set_discontinuity_max_range(cfg->discontinuity_max_range);
    config.set_discontinuity_normal_angle_diff(cfg->discontinuity_normal_angle_diff);
    config.set_discontinuity_z_diff(cfg->discontinuity_z_diff);
    config.set_discontinuity_z_ratio(cfg->discontinuity_z_ratio);
    const eigenmath::Pose3f pose(
        eigenmath::Quaternionf(robot_q_xyzw[3], robot_q_xyzw[0], robot_q_xyzw[1], robot_q_xyzw[2]),
        Vector3f::Zero());
    Region::InputData input(pose, ConstView(xyz, rows, cols), ConstView(normals, rows, cols));
    CloudBuffer<int> lab(rows, cols);
    std::memcpy(lab.Data(), labels, sizeof(int) * static_cast<size_t>(rows) * cols);
    const std::vector<int> seed_vec(seeds, seeds + n_seeds);
    std::vector<Region> regions;
    mobility::SegmentRegions<Region>(input, seed_vec, config, &regions, &lab, id_offset);
    std::memcpy(labels, lab.Data(), sizeof(int) * static_cast<size_t>(rows) * cols);
    for (const Region& r : regions) {
      AppendInliers(res, r.inlier_indices(), r.label_id());
      res->seed_indices.push_back(r.seed_point_index());
      const auto& pl = r.plane();
      res->planes.insert(res->planes.end(), {pl.normal().x(), pl.normal().y(), pl.normal().z(), pl.offset()});
      res->centroids.insert(res->centroids.end(), {r.centroid().x(), r.centroid().y(), r.centroid().z()});
      res->areas.push_back(r.area());
      mobility::PlaneEstimatorProto ep;
      r.plane_estimator().ToProto(&ep);
      for (int i = 0; i < 6; ++i) res->estimators.push_back(ep.covariance_accumulator(i));
      for (int i = 0; i < 3; ++i) res->estimators.push_back(ep.cumulative_centroid(i));
      res->estimators.push_back(ep.cumulative_weights());
      res->boundary_indices.insert(res->boundary_indices.end(), r.boundary_indices().begin(),
                                   r.boundary_indices().end());
      res->boundary_offsets.push_back(static_cast<int>(res->boundary_indices.size()));
      std::vector<int> disc(r.discontinuous_boundary_indices().begin(), r.discontinuous_boundary_indices().end());
      std::sort(disc.begin(), disc.end());
      res->disc_indices.insert(res->disc_indices.end(), disc.begin(), disc.end());
      res->disc_offsets.push_back(static_cast<int>(res->disc_indices.size()));
      const auto& hull = r.projected_boundary_points();
      for (int i = 0; i < hull.Size(); ++i) {
        const Vector3f& h = hull.AtUnsafe(i);
        res->hull_points.insert(res->hull_points.end(), {h.x(), h.y(), h.z()});
      }
      res->hull_offsets.push_back(static_cast<int>(res->hull_points.size() / 3));
    }
  });
}

int oracle_segment_cluster(const float* xyz, int rows, int cols, const int* seeds, int n_seeds,
                           const oracle_cluster_config* cfg, int id_offset, int* labels, oracle_regions** out) {
  auto* res = new oracle_regions;
  *out = res;
  return Guarded([&] {
    using Region = mobility::ClusterRegion<>;
    mobility::ClusterRegionConfigProto config;
    config.set_min_region_inliers(cfg->min_region_inliers);
    config.set_squared_distance_threshold(cfg->squared_distance_threshold);
    config.set_half_search_window(cfg->half_search_window);
    auto points = PointsView(xyz, rows, cols);
    CloudBuffer<int> lab(rows, cols);
    std::memcpy(lab.Data(), labels, sizeof(int) * static_cast<size_t>(rows) * cols);
    const std::vector<int> seed_vec(seeds, seeds + n_seeds);
    std::vector<Region> regions;
    mobility::SegmentRegions<Region>(static_cast<const Cloud<Vector3f>&>(points), seed_vec, config, &regions,
                                     &lab, id_offset);
    std::memcpy(labels, lab.Data(), sizeof(int) * static_cast<size_t>(rows) * cols);
    for (const Region& r : regions) AppendInliers(res, r.inlier_indices(), r.label_id());
  });
}

int oracle_mean_shift(const float* xyz, int rows, int cols, const oracle_cluster_config* cfg, int iterations,
                      int id_offset, int sort_mode, int* labels, oracle_regions** out, float* modes_out,
                      int modes_cap, int* n_modes) {
  auto* res = new oracle_regions;
  *out = res;
  if (n_modes) *n_modes = 0;  // the reference keeps its mode records internal
  if (sort_mode != 0) return 2;  // patch P2 (stable sort) exists only in the port
  (void)modes_out;
  (void)modes_cap;
  return Guarded([&] {
    mobility::ClusterRegionConfigProto config;
    config.set_min_region_inliers(cfg->min_region_inliers);
    config.set_squared_distance_threshold(cfg->squared_distance_threshold);
    config.set_half_search_window(cfg->half_search_window);
    auto points = PointsView(xyz, rows, cols);
    CloudView<int> lab(labels, rows, cols);
    std::vector<mobility::ClusterRegion<>> regions;
    mobility::SlidingMeanShift(static_cast<const Cloud<Vector3f>&>(points), config, iterations, id_offset,
                               static_cast<Cloud<int>*>(&lab), &regions);
    for (const auto& r : regions) {
      AppendInliers(res, r.inlier_indices(), r.label_id());
      const Vector3f s = r.Seed().value();
      res->seed_points.insert(res->seed_points.end(), {s.x(), s.y(), s.z()});
    }
  });
}

int oracle_seeds_plane_support(const float* xyz, const float* normals, int rows, int cols, int neighborhood_size,
                               float max_plane_distance, int min_support, int* seeds_out, int cap) {
  int n = -1;
  Guarded([&] {
    auto p = PointsView(xyz, rows, cols);
    auto nr = PointsView(normals, rows, cols);
    const std::vector<int> s = mobility::FindSeedPointsFromPlaneSupport(
        static_cast<const Cloud<Vector3f>&>(p), static_cast<const Cloud<Vector3f>&>(nr), neighborhood_size,
        max_plane_distance, min_support);
    n = static_cast<int>(s.size());
    for (int i = 0; i < n && i < cap; ++i) seeds_out[i] = s[i];
  });
  return n;
}

int oracle_seeds_average_normals(const float* xyz, const float* normals, int rows, int cols, int neighborhood_size,
                                 int min_num_valid_normals, float min_avg_normal_length, int* seeds_out, int cap) {
  int n = -1;
  Guarded([&] {
    auto p = PointsView(xyz, rows, cols);
    auto nr = PointsView(normals, rows, cols);
    const std::vector<int> s = mobility::FindSeedPointsFromAverageNormals(
        static_cast<const Cloud<Vector3f>&>(p), static_cast<const Cloud<Vector3f>&>(nr), neighborhood_size,
        min_num_valid_normals, min_avg_normal_length);
    n = static_cast<int>(s.size());
    for (int i = 0; i < n && i < cap; ++i) seeds_out[i] = s[i];
  });
  return n;
}

int oracle_seeds_last_planar_regions(const float* xyz, const float* normals, int n_points, int n_prev,
                                     const int* prev_inlier_counts, const float* prev_centroids,
                                     const float* prev_normals, const float q_xyzw[4], const float t[3],
                                     float max_distance, float max_normal_difference_angle, int* seeds_out) {
  auto p = PointsView(xyz, 1, n_points);
  auto nr = PointsView(normals, 1, n_points);
  std::vector<mobility::PlanarRegion<>> prev(n_prev, mobility::PlanarRegion<>(mobility::PlanarRegionConfigProto()));
  for (int i = 0; i < n_prev; ++i) {
    prev[i].MutableInlierIndicesForTestOnly()->assign(prev_inlier_counts[i], 0);
    *prev[i].mutable_centroid() = Vector3f(prev_centroids[3 * i], prev_centroids[3 * i + 1], prev_centroids[3 * i + 2]);
    *prev[i].mutable_plane() = eigenmath::Plane3f(
        Vector3f(prev_normals[3 * i], prev_normals[3 * i + 1], prev_normals[3 * i + 2]), Vector3f(0.0f, 0.0f, 0.0f));
  }
  const eigenmath::Pose3f pose(eigenmath::Quaternionf(q_xyzw[3], q_xyzw[0], q_xyzw[1], q_xyzw[2]),
                               Vector3f(t[0], t[1], t[2]));
  std::vector<int> seeds;
  mobility::FindSeedPointsFromLastPlanarRegions(static_cast<const Cloud<Vector3f>&>(p),
                                                static_cast<const Cloud<Vector3f>&>(nr), prev, pose, max_distance,
                                                max_normal_difference_angle, &seeds);
  for (size_t i = 0; i < seeds.size(); ++i) seeds_out[i] = seeds[i];
  return static_cast<int>(seeds.size());
}

int oracle_regions_count(const oracle_regions* r) { return static_cast<int>(r->label_ids.size()); }
const int* oracle_regions_inlier_offsets(const oracle_regions* r) { return r->inlier_offsets.data(); }
const int* oracle_regions_inlier_indices(const oracle_regions* r) { return r->inlier_indices.data(); }
const int* oracle_regions_label_ids(const oracle_regions* r) { return r->label_ids.data(); }
const int* oracle_regions_seed_indices(const oracle_regions* r) { return r->seed_indices.data(); }
const float* oracle_regions_planes(const oracle_regions* r) { return r->planes.data(); }
const float* oracle_regions_centroids(const oracle_regions* r) { return r->centroids.data(); }
const float* oracle_regions_areas(const oracle_regions* r) { return r->areas.data(); }
const float* oracle_regions_estimators(const oracle_regions* r) { return r->estimators.data(); }
const int* oracle_regions_boundary_offsets(const oracle_regions* r) { return r->boundary_offsets.data(); }
const int* oracle_regions_boundary_indices(const oracle_regions* r) { return r->boundary_indices.data(); }
const int* oracle_regions_discontinuous_offsets(const oracle_regions* r) { return r->disc_offsets.data(); }
const int* oracle_regions_discontinuous_indices(const oracle_regions* r) { return r->disc_indices.data(); }
const int* oracle_regions_hull_offsets(const oracle_regions* r) { return r->hull_offsets.data(); }
const float* oracle_regions_hull_points(const oracle_regions* r) { return r->hull_points.data(); }
const float* oracle_regions_seed_points(const oracle_regions* r) { return r->seed_points.data(); }
void oracle_regions_free(oracle_regions* r) { delete r; }

}  // extern "C"
