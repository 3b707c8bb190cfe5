// The reading half of sensor::CompressedPointCloud (sensor/compressed_point_cloud.{h,cc}): the
// form in which a serialized trajectory node carries its point clouds
// (mapping/proto/trajectory_node_data.proto fields 3-5, unpacked by mapping/trajectory_node.cc:50-69).
// Layout of point_data (compressed_point_cloud.cc:30-36,80-101,128-145): blocks of
//   [points in block, block x, block y, block z, point, point, ...]
// with a point = z << 20 | y << 10 | x, 10 bits per coordinate relative to the block origin
// (block coordinate << 10), in units of kPrecision = 1 mm.  A decoded coordinate is the int32 sum
// converted to float and multiplied by 0.001f -- one correctly rounded product, so the decoded
// floats are bit-identical to the reference's.
// The compressing constructor (cc:103-146) is not restated: nothing on the scan-matching path
// writes nodes.
#ifndef CARTOGRAPHER_SENSOR_COMPRESSED_POINT_CLOUD_H_
#define CARTOGRAPHER_SENSOR_COMPRESSED_POINT_CLOUD_H_

#include <cstdint>
#include <vector>

#include "cartographer_compat.h"

namespace cartographer {
namespace sensor {

class CompressedPointCloud {
 public:
  CompressedPointCloud() = default;
  // What CompressedPointCloud(const proto::CompressedPointCloud&) copies out of the message
  // (cc:148-157).
  CompressedPointCloud(const size_t num_points, std::vector<int32_t> point_data)
      : point_data_(std::move(point_data)), num_points_(num_points) {}

  bool empty() const { return num_points_ == 0; }
  size_t size() const { return num_points_; }
  const std::vector<int32_t>& point_data() const { return point_data_; }

  // cc:44-101,169-175.  The reference trusts point_data ("TODO: verify that point_data_ does not
  // contain malformed data"); here a stream that ends before num_points points were read aborts.
  PointCloud Decompress() const {
    constexpr float kPrecision = 0.001f;
    constexpr int kBitsPerCoordinate = 10;
    constexpr int kMask = (1 << kBitsPerCoordinate) - 1;
    std::vector<RangefinderPoint> points;
    points.reserve(num_points_);
    size_t input = 0;
    int32_t remaining_points_in_current_block = 0;
    int32_t block[3] = {0, 0, 0};
    for (size_t remaining = num_points_; remaining > 0; --remaining) {
      if (remaining_points_in_current_block == 0) {
        CSM_CHECK(input + 4 <= point_data_.size());
        remaining_points_in_current_block = point_data_[input++];
        CSM_CHECK(remaining_points_in_current_block > 0);
        for (int i = 0; i < 3; ++i)
          block[i] = static_cast<int32_t>(static_cast<uint32_t>(point_data_[input++]) << kBitsPerCoordinate);
      }
      --remaining_points_in_current_block;
      CSM_CHECK(input < point_data_.size());
      const int32_t point = point_data_[input++];
      points.push_back({Vector3f((block[0] + (point & kMask)) * kPrecision,
                                 (block[1] + ((point >> kBitsPerCoordinate) & kMask)) * kPrecision,
                                 (block[2] + (point >> (2 * kBitsPerCoordinate))) * kPrecision)});
    }
    return PointCloud(std::move(points));
  }

  bool operator==(const CompressedPointCloud& o) const {
    return point_data_ == o.point_data_ && num_points_ == o.num_points_;
  }

 private:
  std::vector<int32_t> point_data_;
  size_t num_points_ = 0;
};

}  // namespace sensor
}  // namespace cartographer

#endif  // CARTOGRAPHER_SENSOR_COMPRESSED_POINT_CLOUD_H_
