// TEST INFRASTRUCTURE ONLY.  Stand-in for absl/container/flat_hash_set.h: the reference only
// inserts, clears and iterates (ps/planar_region.h:303,414,435-437); results are compared as sets.
#ifndef ORACLE_SHIMS_ABSL_FLAT_HASH_SET_H_
#define ORACLE_SHIMS_ABSL_FLAT_HASH_SET_H_
#include <unordered_set>
namespace absl {
template <typename T>
using flat_hash_set = std::unordered_set<T>;
}  // namespace absl
#endif
