// TEST INFRASTRUCTURE (oracle/_ref build only): abseil is not in this image.  The voxel filter
// uses flat_hash_map as a plain key -> value table (operator[], iteration whose order does not
// reach the result: it only sets per-point flags), so std::unordered_map stands in.
#ifndef ORACLE_REF_SHIM_ABSL_FLAT_HASH_MAP_H_
#define ORACLE_REF_SHIM_ABSL_FLAT_HASH_MAP_H_
#include <unordered_map>
namespace absl {
template <typename K, typename V>
using flat_hash_map = std::unordered_map<K, V>;
}  // namespace absl
#endif
