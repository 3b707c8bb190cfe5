// TEST INFRASTRUCTURE (oracle/_ref build only): Ceres is not in this image; common/math.h uses
// ceres::atan2 on plain scalars only along this path.
#ifndef ORACLE_REF_SHIM_CERES_H_
#define ORACLE_REF_SHIM_CERES_H_
#include <cmath>
namespace ceres {
using std::atan2;
using std::sqrt;
using std::sin;
using std::cos;
using std::abs;
}  // namespace ceres
#endif
