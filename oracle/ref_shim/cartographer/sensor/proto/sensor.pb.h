// TEST INFRASTRUCTURE (oracle/_ref build only): stand-in for protoc's output of
// cartographer/sensor/proto/sensor.proto (the messages the point-cloud headers touch).
#ifndef ORACLE_REF_SHIM_SENSOR_PB_H_
#define ORACLE_REF_SHIM_SENSOR_PB_H_
#include "cartographer/transform/proto/transform.pb.h"
namespace cartographer {
namespace sensor {
namespace proto {
class RangefinderPoint { SHIM_MESSAGE(transform::proto::Vector3f, position) };
class TimedRangefinderPoint { SHIM_MESSAGE(transform::proto::Vector3f, position) SHIM_SCALAR(float, time) };
class CompressedPointCloud { SHIM_SCALAR(int32_t, num_points) SHIM_REPEATED(int32_t, point_data) };
class TimedPointCloudData {
  SHIM_SCALAR(int64_t, timestamp)
  SHIM_MESSAGE(transform::proto::Vector3f, origin)
  SHIM_REPEATED(transform::proto::Vector4f, point_data_legacy)
  SHIM_REPEATED(TimedRangefinderPoint, point_data)
  SHIM_REPEATED(float, intensities)
};
class RangeData {
  SHIM_MESSAGE(transform::proto::Vector3f, origin)
  SHIM_REPEATED(transform::proto::Vector3f, returns_legacy)
  SHIM_REPEATED(transform::proto::Vector3f, misses_legacy)
  SHIM_REPEATED(RangefinderPoint, returns)
  SHIM_REPEATED(RangefinderPoint, misses)
};
}  // namespace proto
}  // namespace sensor
}  // namespace cartographer
#endif
