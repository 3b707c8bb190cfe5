// TEST INFRASTRUCTURE (oracle/_ref build only): abseil is not in this image.
#ifndef ORACLE_REF_SHIM_ABSL_OPTIONAL_H_
#define ORACLE_REF_SHIM_ABSL_OPTIONAL_H_
#include <optional>
namespace absl {
using std::optional;
using std::nullopt;
using std::make_optional;
}  // namespace absl
#endif
