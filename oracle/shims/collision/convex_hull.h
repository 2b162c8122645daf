// TEST INFRASTRUCTURE ONLY.  Stand-in for collision/convex_hull.h (x_edr_mobility_collision
// 1.0.0, WORKSPACE:86-93, not vendored).  Restated as Andrew's monotone chain in double:
// counter-clockwise, duplicates and collinear points dropped.  Only the number of hull
// points gates labels (ps/planar_region.h:217); vertex order of the real library is unknown.
#ifndef ORACLE_SHIMS_COLLISION_CONVEX_HULL_H_
#define ORACLE_SHIMS_COLLISION_CONVEX_HULL_H_
#include <algorithm>
#include <vector>
#include "eigenmath/types.h"
namespace mobility {
namespace collision {
class ConvexHull {
 public:
  explicit ConvexHull(const std::vector<eigenmath::Vector2d>& in) {
    std::vector<eigenmath::Vector2d> p = in;
    std::sort(p.begin(), p.end(), [](const eigenmath::Vector2d& a, const eigenmath::Vector2d& b) {
      return a.x() < b.x() || (a.x() == b.x() && a.y() < b.y());
    });
    p.erase(std::unique(p.begin(), p.end(),
                        [](const eigenmath::Vector2d& a, const eigenmath::Vector2d& b) {
                          return a.x() == b.x() && a.y() == b.y();
                        }),
            p.end());
    const int n = static_cast<int>(p.size());
    if (n < 3) {
      points_ = p;
      return;
    }
    std::vector<eigenmath::Vector2d> h(2 * n);
    int k = 0;
    for (int i = 0; i < n; ++i) {
      while (k >= 2 && Cross(h[k - 2], h[k - 1], p[i]) <= 0) --k;
      h[k++] = p[i];
    }
    for (int i = n - 2, t = k + 1; i >= 0; --i) {
      while (k >= t && Cross(h[k - 2], h[k - 1], p[i]) <= 0) --k;
      h[k++] = p[i];
    }
    h.resize(k - 1);
    points_ = h;
  }
  const std::vector<eigenmath::Vector2d>& GetPoints() const { return points_; }
 private:
  static double Cross(const eigenmath::Vector2d& o, const eigenmath::Vector2d& a, const eigenmath::Vector2d& b) {
    return (a.x() - o.x()) * (b.y() - o.y()) - (a.y() - o.y()) * (b.x() - o.x());
  }
  std::vector<eigenmath::Vector2d> points_;
};
}  // namespace collision
}  // namespace mobility
#endif
