// C++ parity test through the host-side mirror (pcseg_host.h) of the reference API.
// Reads like ps/clustering_test.cc:115-151 (known-answer test of
// SegmentRegions<ClusterRegion>): 10x10 column-major cloud, seeds stored reversed so
// that they are consumed in ascending index order, default config, id offset 1; then the
// same scene through SegmentRegions<PlanarRegion> / ComputeNormalsOrganized /
// SlidingMeanShift as smoke checks of the remaining entry points.
#include <cmath>
#include <cstdio>
#include <vector>

#include "pcseg_host.h"

namespace B = mobility::b200;

using B::CloudRef;
using B::Float3;

namespace {
constexpr int kRows = 10, kCols = 10;

// Scene of SURVEY.md Appendix C.1 (restated from the reference test's description).
std::vector<Float3> PointCloudData() {
  std::vector<Float3> p(kRows * kCols);
  for (int c = 0; c < kCols; ++c) {
    for (int r = 0; r < kRows; ++r) {
      Float3 v;
      if (c < 4) {
        if (r < 4) v = {0.05f * c + 2, 0.05f * r + 2, 2};
        else if (r > 4) v = {0.05f * c, 0.05f * r + 2, 0};
        else v = {0.05f * c, 0.05f * r + 2, 1};
      } else if (c == 4) {
        v = {0.1f * c, 0, 0.1f * r};
      } else if (r + c == 9 || r + c == 10) {
        v = {0.05f * c + 5, -0.05f * r - 2, 2 - 0.05f * r};
      } else if (r == 8 && c == 8) {
        v = {0.05f * c + 2, -0.05f * r - 2, 4};
      } else if ((r == 6 || r == 7) && c == 8) {
        v = {6, 8, 2};
      } else {
        v = {0.05f * c + 2, -0.05f * r - 2, 2 - 0.05f * r};
      }
      p[c * kRows + r] = v;
    }
  }
  return p;
}

const int kExpected[kRows][kCols] = {  // ps/clustering_test.cc:127-137, [row][col]
    {1, 1, 1, 1, 3, 4, 4, 4, 4, 5},   {1, 1, 1, 1, 3, 4, 4, 4, 5, 5},   {1, 1, 1, 1, 3, 4, 4, 5, 5, 6},
    {1, 1, 1, 1, 3, 4, 5, 5, 6, 6},   {-1, -1, -1, -1, 3, 5, 5, 6, 6, 6}, {2, 2, 2, 2, 3, 5, 6, 6, 6, 6},
    {2, 2, 2, 2, 3, 6, 6, 6, -1, 6},  {2, 2, 2, 2, 3, 6, 6, 6, -1, 6},  {2, 2, 2, 2, 3, 6, 6, 6, -1, 6},
    {2, 2, 2, 2, 3, 6, 6, 6, 6, 6}};

int g_failures = 0;
#define EXPECT(c)                                                   \
  do {                                                              \
    if (!(c)) {                                                     \
      std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #c);    \
      ++g_failures;                                                 \
    }                                                               \
  } while (0)
}  // namespace

int main() {
  B::Context ctx(0);
  std::vector<Float3> data = PointCloudData();
  CloudRef<Float3> points(data.data(), kRows, kCols);
  std::vector<int> seeds(kRows * kCols);
  for (int i = 0; i < kRows * kCols; ++i) seeds[kRows * kCols - 1 - i] = i;

  // --- ps/clustering_test.cc:115-151 ---
  std::vector<int> label_store = B::PrepareUnlabeledLabels(ctx, points);
  CloudRef<int> labels(label_store.data(), kRows, kCols);
  std::vector<B::ClusterRegion> regions;
  B::SegmentRegions(ctx, points, seeds, B::ClusterRegionConfig(), &regions, &labels, 1);
  for (int r = 0; r < kRows; ++r)
    for (int c = 0; c < kCols; ++c) EXPECT(label_store[labels.LinearizeIndex(r, c)] == kExpected[r][c]);
  const size_t sizes[] = {16, 20, 10, 10, 10, 27};
  EXPECT(regions.size() == 6);
  for (size_t i = 0; i < regions.size() && i < 6; ++i) {
    EXPECT(regions[i].inlier_indices().size() == sizes[i]);
    EXPECT(regions[i].label_id() == (int)i + 1);
  }

  // --- normals + planes on a tilted 64x64 plane with a step ---
  const int R = 64, C = 64;
  std::vector<Float3> plane(R * C), normals(R * C);
  for (int c = 0; c < C; ++c)
    for (int r = 0; r < R; ++r)
      plane[c * R + r] = {0.03f * c, 0.03f * r, (c < 32 ? 2.0f : 2.6f) + 0.002f * r};
  CloudRef<Float3> pc(plane.data(), R, C), nc(normals.data(), R, C);
  const double sensor[3] = {0, 0, 0};
  B::ComputeNormalsParams np;
  np.min_neighbor_distance = 0.02f;
  B::ComputeNormalsOrganized(ctx, sensor, pc, &nc, np);
  int finite = 0;
  for (const Float3& n : normals) finite += std::isfinite(n.x) ? 1 : 0;
  EXPECT(finite > R * C / 2);
  EXPECT(normals[10 * R + 10].z < -0.99f);  // oriented towards the sensor at the origin
  std::vector<int> lab2 = B::PrepareUnlabeledLabels(ctx, pc);
  CloudRef<int> labels2(lab2.data(), R, C);
  std::vector<int> seeds2 = B::FindSeedPointsFromPlaneSupport(ctx, pc, nc);
  EXPECT(!seeds2.empty());
  std::vector<B::PlanarRegion> planes;
  B::PlanarRegion::InputData<CloudRef<Float3>, CloudRef<Float3>> in{{0, 0, 0, 1}, pc, nc};
  B::SegmentRegions(ctx, in, seeds2, B::PlanarRegionConfig(), &planes, &labels2, 0);
  EXPECT(planes.size() == 2);
  for (const auto& p : planes) {
    EXPECT(p.inlier_indices().size() > 1500);
    EXPECT(std::fabs(std::fabs(p.plane()[2]) - 1.0f) < 1e-2f);
    EXPECT(p.area() > 1.0f);
    EXPECT(!p.boundary_indices().empty());
    EXPECT(p.projected_boundary_points().size() >= 3);
  }

  // --- SlidingMeanShift on the same sheet (labels reset) ---
  std::vector<int> lab3 = B::PrepareUnlabeledLabels(ctx, pc);
  CloudRef<int> labels3(lab3.data(), R, C);
  std::vector<B::ClusterRegion> ms;
  B::SlidingMeanShift(ctx, pc, B::ClusterRegionConfig(), 3, 0, &labels3, &ms);
  EXPECT(!ms.empty());
  for (const auto& m : ms) EXPECT(m.Seed().has_value() && m.inlier_indices().size() >= 7);

  // --- PlaneEstimator: points of the plane z = 1 + 0.1 x (ps/plane_estimator_test.cc:40-74 style) ---
  {
    B::PlaneEstimator est(ctx);
    for (int i = 0; i < 5; ++i)
      for (int j = 0; j < 5; ++j) est.AddPoint(Float3{0.1f * i, 0.1f * j, 1.0f + 0.01f * i});
    est.SetNormalOrientation(Float3{0, 0, 1});
    EXPECT(est.PlaneValid());
    const float* pl = est.Plane();
    EXPECT(pl[2] > 0.99f && std::fabs(pl[0] + 0.1f * pl[2]) < 1e-4f && std::fabs(pl[1]) < 1e-4f);
    const Float3 c = est.Centroid();
    EXPECT(std::fabs(c.x - 0.2f) < 1e-5f && std::fabs(c.y - 0.2f) < 1e-5f && std::fabs(c.z - 1.02f) < 1e-5f);
    EXPECT(est.CumulativeWeight() == 25.0f);
    // the same fit through the batched CSR entry point
    std::vector<Float3> pts;
    for (int i = 0; i < 5; ++i)
      for (int j = 0; j < 5; ++j) pts.push_back(Float3{0.1f * i, 0.1f * j, 1.0f + 0.01f * i});
    CloudRef<Float3> pcl(pts.data(), 5, 5);
    std::vector<int> off = {0, 25}, idx(25);
    for (int i = 0; i < 25; ++i) idx[i] = i;
    std::vector<float> q;
    B::FitPlanes(ctx, pcl, off, idx, &q);
    EXPECT(q.size() == 10 && q[0] == 1.0f && std::fabs(std::fabs(q[7]) - std::fabs(pl[2])) < 1e-6f);
  }
  // --- TransformInPlace + DetectedObject geometry gathers ---
  {
    std::vector<Float3> tp(plane), tn(normals);
    CloudRef<Float3> tpc(tp.data(), R, C), tnc(tn.data(), R, C);
    const double q[4] = {0, 0, 0, 1}, tr[3] = {1.0, -2.0, 0.5};
    B::TransformInPlace(ctx, q, tr, &tpc, &tnc);
    EXPECT(std::fabs(tp[5].x - (plane[5].x + 1.0f)) < 1e-6f && std::fabs(tp[5].y - (plane[5].y - 2.0f)) < 1e-6f);
    if (!planes.empty()) {
      const B::DetectedObjectGeometry g = B::CreatePlanarDetectedObjectGeometry(ctx, pc, planes[0]);
      EXPECT(g.points.size() == planes[0].inlier_indices().size());
      const int i0 = planes[0].inlier_indices()[0];
      EXPECT(g.points[0].x == plane[i0].x && g.points[0].z == plane[i0].z);
      EXPECT(g.discontinuous_positions.size() == planes[0].discontinuous_boundary_indices().size());
    }
    if (!regions.empty()) {
      const B::DetectedObjectGeometry g = B::CreateClusterDetectedObjectGeometry(ctx, points, regions[0]);
      EXPECT(g.points.size() == regions[0].inlier_indices().size());
    }
  }
  // --- the other two seed finders (shape checks; bit parity is pinned by tests/test_estimator_seeds.py) ---
  {
    std::vector<int> s_avg = B::FindSeedPointsFromAverageNormals(ctx, pc, nc);
    EXPECT(!s_avg.empty());
    std::vector<int> s_last;
    const float q[4] = {0, 0, 0, 1}, tr[3] = {0, 0, 0};
    B::FindSeedPointsFromLastPlanarRegions(ctx, pc, nc, planes, q, tr, 0.5f, 0.3f, &s_last);
    EXPECT(s_last.size() <= planes.size());
  }

  // --- unorganized clustering (extension): two blobs 2 m apart and a lone point ---
  {
    std::vector<float> u;
    for (int k = 0; k < 40; ++k) { u.push_back(0.01f * k); u.push_back(0.0f); u.push_back(1.0f); }
    for (int k = 0; k < 25; ++k) { u.push_back(2.0f + 0.01f * k); u.push_back(0.5f); u.push_back(1.0f); }
    u.push_back(9.0f); u.push_back(9.0f); u.push_back(9.0f);
    std::vector<int> ul(66, -1);
    std::vector<B::ClusterRegion> ur;
    B::ClusterUnorganized(ctx, u.data(), 66, 0.1f, 5, 3, ul.data(), &ur);
    EXPECT(ur.size() == 2);
    if (ur.size() == 2) {
      EXPECT(ur[0].label_id() == 3 && ur[1].label_id() == 4);
      EXPECT(ur[0].inlier_indices().size() == 40 && ur[1].inlier_indices().size() == 25);
      EXPECT(ur[0].Centroid() && std::fabs(ur[0].Centroid()->x - 0.195f) < 1e-5f);
      EXPECT(ur[1].ExtentMin() && ur[1].ExtentMin()->x == 2.0f && ur[1].ExtentMax()->y == 0.5f);
    }
    EXPECT(ul[0] == 3 && ul[64] == 4 && ul[65] == -1);
  }

  if (g_failures == 0) std::printf("PASS clustering_host_test: %zu clusters, %zu planes, %zu mean-shift regions\n",
                                   regions.size(), planes.size(), ms.size());
  return g_failures == 0 ? 0 : 1;
}
