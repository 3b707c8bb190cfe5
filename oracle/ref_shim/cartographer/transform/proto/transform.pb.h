// TEST INFRASTRUCTURE (oracle/_ref build only): stand-in for protoc's output of
// cartographer/transform/proto/transform.proto.
#ifndef ORACLE_REF_SHIM_TRANSFORM_PB_H_
#define ORACLE_REF_SHIM_TRANSFORM_PB_H_
#include "shim_proto.h"
namespace cartographer {
namespace transform {
namespace proto {
class Vector2d { SHIM_SCALAR(double, x) SHIM_SCALAR(double, y) };
class Vector2f { SHIM_SCALAR(float, x) SHIM_SCALAR(float, y) };
class Vector3d { SHIM_SCALAR(double, x) SHIM_SCALAR(double, y) SHIM_SCALAR(double, z) };
class Vector3f { SHIM_SCALAR(float, x) SHIM_SCALAR(float, y) SHIM_SCALAR(float, z) };
class Vector4f { SHIM_SCALAR(float, x) SHIM_SCALAR(float, y) SHIM_SCALAR(float, z) SHIM_SCALAR(float, t) };
class Quaterniond { SHIM_SCALAR(double, x) SHIM_SCALAR(double, y) SHIM_SCALAR(double, z) SHIM_SCALAR(double, w) };
class Quaternionf { SHIM_SCALAR(float, x) SHIM_SCALAR(float, y) SHIM_SCALAR(float, z) SHIM_SCALAR(float, w) };
class Rigid2d { SHIM_MESSAGE(Vector2d, translation) SHIM_SCALAR(double, rotation) };
class Rigid2f { SHIM_MESSAGE(Vector2f, translation) SHIM_SCALAR(float, rotation) };
class Rigid3d { SHIM_MESSAGE(Vector3d, translation) SHIM_MESSAGE(Quaterniond, rotation) };
class Rigid3f { SHIM_MESSAGE(Vector3f, translation) SHIM_MESSAGE(Quaternionf, rotation) };
}  // namespace proto
}  // namespace transform
}  // namespace cartographer
#endif
