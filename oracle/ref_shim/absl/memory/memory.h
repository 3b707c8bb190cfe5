// TEST INFRASTRUCTURE (oracle/_ref build only): abseil is not in this image.
#ifndef ORACLE_REF_SHIM_ABSL_MEMORY_H_
#define ORACLE_REF_SHIM_ABSL_MEMORY_H_
#include <memory>
namespace absl {
using std::make_unique;
template <typename T>
std::unique_ptr<T> WrapUnique(T* p) {
  return std::unique_ptr<T>(p);
}
}  // namespace absl
#endif
