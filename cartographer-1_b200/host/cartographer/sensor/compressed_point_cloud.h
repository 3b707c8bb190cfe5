// sensor::CompressedPointCloud (sensor/compressed_point_cloud.{h,cc}): the
// form in which a serialized trajectory node carries its point clouds
// (mapping/proto/trajectory_node_data.proto fields 3-5, unpacked by mapping/trajectory_node.cc:50-69).
// Layout of point_data (compressed_point_cloud.cc:30-36,80-101,128-145): blocks of
//   [points in block, block x, block y, block z, point, point, ...]
// with a point = z << 20 | y << 10 | x, 10 bits per coordinate relative to the block origin
// (block coordinate << 10), in units of kPrecision = 1 mm.  A decoded coordinate is the int32 sum
// converted to float and multiplied by 0.001f -- one correctly rounded product, so the decoded
// floats are bit-identical to the reference's.
// The compressing constructor (cc:103-146) groups the millimetre raster into blocks through a
// HybridGridBase<std::vector<RasterPoint>> and writes the blocks in that grid's iteration order:
// meta cells of 64^3 blocks, then 8^3 bricks, then the 8^3 blocks of a brick, each level in flat
// z-major order (mapping/3d/hybrid_grid.h:40-52,255-345), the grid having doubled from 128 until the
// outermost block fitted.  Here the blocks are sorted by that key, so point_data() is word for word
// what the reference serializes.
#ifndef CARTOGRAPHER_SENSOR_COMPRESSED_POINT_CLOUD_H_
#define CARTOGRAPHER_SENSOR_COMPRESSED_POINT_CLOUD_H_

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <unordered_map>
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

  explicit CompressedPointCloud(const PointCloud& point_cloud) : num_points_(point_cloud.size()) {
    constexpr float kPrecision = 0.001f;
    constexpr int kBitsPerCoordinate = 10, kCoordinateMask = (1 << kBitsPerCoordinate) - 1;
    constexpr int kMaxBitsPerDirection = 23;
    struct Block {
      int32_t coordinate[3];
      std::vector<int32_t> points;  // packed, in point order
    };
    std::vector<Block> blocks;
    std::unordered_map<uint64_t, size_t> block_of;
    int grid_size = 128;  // hybrid_grid.h:255-260: 2 x 64, doubled by Grow()
    for (const RangefinderPoint& point : point_cloud) {
      const Vector3f& p = point.position;
      const float reach = std::max(std::abs(p.x()), std::max(std::abs(p.y()), std::abs(p.z())));
      CSM_CHECK(reach / kPrecision < (1 << kMaxBitsPerDirection));  // "Point out of bounds"
      int32_t raster[3], block[3];
      const float v[3] = {p.x(), p.y(), p.z()};
      for (int i = 0; i < 3; ++i) {
        raster[i] = common::RoundToInt(v[i] / kPrecision);
        block[i] = raster[i] >> kBitsPerCoordinate;
        raster[i] &= kCoordinateMask;
        while (block[i] < -(grid_size >> 1) || block[i] >= (grid_size >> 1)) grid_size <<= 1;
      }
      const uint64_t key = (static_cast<uint64_t>(static_cast<uint32_t>(block[0] + (1 << 20))) << 42) |
                           (static_cast<uint64_t>(static_cast<uint32_t>(block[1] + (1 << 20))) << 21) |
                           static_cast<uint64_t>(static_cast<uint32_t>(block[2] + (1 << 20)));
      auto it = block_of.find(key);
      if (it == block_of.end()) {
        it = block_of.emplace(key, blocks.size()).first;
        blocks.push_back(Block{{block[0], block[1], block[2]}, {}});
      }
      blocks[it->second].points.push_back((((raster[2] << kBitsPerCoordinate) + raster[1]) << kBitsPerCoordinate) +
                                          raster[0]);
    }
    // iteration order of the reference's block grid at its final size
    int meta_bits = 0;
    while ((64 << meta_bits) < grid_size) ++meta_bits;
    auto order_key = [&](const Block& b) {
      const uint32_t u[3] = {static_cast<uint32_t>(b.coordinate[0] + (grid_size >> 1)),
                             static_cast<uint32_t>(b.coordinate[1] + (grid_size >> 1)),
                             static_cast<uint32_t>(b.coordinate[2] + (grid_size >> 1))};
      auto flat = [](uint32_t x, uint32_t y, uint32_t z, int bits) {
        return static_cast<uint64_t>((((z << bits) + y) << bits) + x);
      };
      const uint64_t meta = flat(u[0] >> 6, u[1] >> 6, u[2] >> 6, meta_bits);
      const uint64_t brick = flat((u[0] >> 3) & 7, (u[1] >> 3) & 7, (u[2] >> 3) & 7, 3);
      const uint64_t cell = flat(u[0] & 7, u[1] & 7, u[2] & 7, 3);
      return (meta << 18) | (brick << 9) | cell;
    };
    std::sort(blocks.begin(), blocks.end(),
              [&](const Block& a, const Block& b) { return order_key(a) < order_key(b); });
    point_data_.reserve(4 * blocks.size() + point_cloud.size());
    for (const Block& b : blocks) {
      point_data_.push_back(static_cast<int32_t>(b.points.size()));
      point_data_.push_back(b.coordinate[0]);
      point_data_.push_back(b.coordinate[1]);
      point_data_.push_back(b.coordinate[2]);
      point_data_.insert(point_data_.end(), b.points.begin(), b.points.end());
    }
  }

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
