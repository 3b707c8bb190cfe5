#include "cartographer/sensor/internal/voxel_filter.h"

#include <cmath>
#include <cstdint>
#include <random>
#include <unordered_map>
#include <utility>

#include "cartographer/common/lua_parameter_dictionary.h"

namespace cartographer {
namespace sensor {
namespace {

// cc:80-87.  The three rounded indices are converted to uint64 (negative ones sign-extend) and
// combined with wrapping arithmetic; two voxels share a key exactly when they do in the reference.
using VoxelKeyType = uint64_t;

VoxelKeyType GetVoxelCellIndex(const Vector3f& point, const float resolution) {
  const uint64_t x = static_cast<uint64_t>(static_cast<int64_t>(common::RoundToInt(point.x() / resolution)));
  const uint64_t y = static_cast<uint64_t>(static_cast<int64_t>(common::RoundToInt(point.y() / resolution)));
  const uint64_t z = static_cast<uint64_t>(static_cast<int64_t>(common::RoundToInt(point.z() / resolution)));
  return (x << 42) + (y << 21) + z;
}

// Per voxel: (points seen, index of the point currently kept).  The reservoir walk shared by the
// two filters (cc:96-112,144-158; en.wikipedia.org/wiki/Reservoir_sampling): the k-th point of a
// voxel replaces the kept one with probability 1 / k.
using VoxelTable = std::unordered_map<VoxelKeyType, std::pair<int, int>>;

VoxelTable ReservoirWalk(const std::vector<RangefinderPoint>& points, const float resolution,
                         std::vector<VoxelKeyType>* keys) {
  std::minstd_rand0 generator;
  VoxelTable voxel_count_and_point_index;
  voxel_count_and_point_index.reserve(points.size());
  if (keys != nullptr) keys->resize(points.size());
  for (size_t i = 0; i < points.size(); i++) {
    const VoxelKeyType key = GetVoxelCellIndex(points[i].position, resolution);
    if (keys != nullptr) (*keys)[i] = key;
    std::pair<int, int>& voxel = voxel_count_and_point_index[key];
    voxel.first++;
    if (voxel.first == 1) {
      voxel.second = static_cast<int>(i);
    } else {
      std::uniform_int_distribution<> distribution(1, voxel.first);
      if (distribution(generator) == voxel.first) voxel.second = static_cast<int>(i);
    }
  }
  return voxel_count_and_point_index;
}

// cc:32-36; Eigen's norm() of a 3-vector: sqrt((x*x + y*y) + z*z) in float.
PointCloud FilterByMaxRange(const PointCloud& point_cloud, const float max_range) {
  std::vector<RangefinderPoint> kept;
  for (const RangefinderPoint& point : point_cloud) {
    const Vector3f& p = point.position;
    if (std::sqrt(p.x() * p.x() + p.y() * p.y() + p.z() * p.z()) <= max_range) kept.push_back(point);
  }
  return PointCloud(std::move(kept));
}

// cc:38-76
PointCloud AdaptivelyVoxelFiltered(const proto::AdaptiveVoxelFilterOptions& options, const PointCloud& point_cloud) {
  if (point_cloud.size() <= options.min_num_points()) return point_cloud;  // already sparse enough
  PointCloud result = VoxelFilter(point_cloud, options.max_length());
  if (result.size() >= options.min_num_points()) return result;
  // Halve the edge until enough points remain (giving up at 1e-2 of max_length), then bisect until
  // the edge is within 10 % of the largest one known to keep enough.
  for (float high_length = options.max_length(); high_length > 1e-2f * options.max_length(); high_length /= 2.f) {
    float low_length = high_length / 2.f;
    result = VoxelFilter(point_cloud, low_length);
    if (result.size() >= options.min_num_points()) {
      while ((high_length - low_length) / low_length > 1e-1f) {
        const float mid_length = (low_length + high_length) / 2.f;
        const PointCloud candidate = VoxelFilter(point_cloud, mid_length);
        if (candidate.size() >= options.min_num_points()) {
          low_length = mid_length;
          result = candidate;
        } else {
          high_length = mid_length;
        }
      }
      return result;
    }
  }
  return result;
}

PointCloud Select(const PointCloud& point_cloud, const std::vector<bool>& points_used) {
  std::vector<RangefinderPoint> filtered_points;
  for (size_t i = 0; i < point_cloud.size(); i++)
    if (points_used[i]) filtered_points.push_back(point_cloud[i]);
  return PointCloud(std::move(filtered_points));
}

}  // namespace

std::vector<bool> VoxelFilterIndices(const std::vector<RangefinderPoint>& points, const float resolution) {
  const VoxelTable table = ReservoirWalk(points, resolution, nullptr);
  std::vector<bool> points_used(points.size(), false);
  for (const auto& voxel_and_index : table) points_used[voxel_and_index.second.second] = true;
  return points_used;
}

// cc:89-137,165-204: a point is kept when its voxel (edge `rad`) holds fewer than
// int(max voxel population * voxeledgeratio) points.  (The reference runs the reservoir walk here
// too and ignores its choices; only the counts matter.)
PointCloud VoxelFilterEdge(const PointCloud& point_cloud, const float rad, const float voxeledgeratio) {
  std::vector<VoxelKeyType> keys;
  const VoxelTable table = ReservoirWalk(point_cloud.points(), rad, &keys);
  int maxneighbors = 0;
  for (const auto& voxel : table) maxneighbors = std::max(maxneighbors, voxel.second.first);
  const int limit = static_cast<int>(maxneighbors * voxeledgeratio);
  std::vector<bool> points_used(point_cloud.size(), false);
  for (size_t i = 0; i < point_cloud.size(); i++) points_used[i] = table.at(keys[i]).first < limit;
  return Select(point_cloud, points_used);
}

std::vector<RangefinderPoint> VoxelFilter(const std::vector<RangefinderPoint>& points, const float resolution) {
  const std::vector<bool> points_used = VoxelFilterIndices(points, resolution);
  std::vector<RangefinderPoint> results;
  for (size_t i = 0; i < points.size(); i++)
    if (points_used[i]) results.push_back(points[i]);
  return results;
}

PointCloud VoxelFilter(const PointCloud& point_cloud, const float resolution) {
  return Select(point_cloud, VoxelFilterIndices(point_cloud.points(), resolution));
}

// cc:253-262
proto::AdaptiveVoxelFilterOptions CreateAdaptiveVoxelFilterOptions(
    common::LuaParameterDictionary* const parameter_dictionary) {
  proto::AdaptiveVoxelFilterOptions options;
  options.set_max_length(static_cast<float>(parameter_dictionary->GetDouble("max_length")));
  options.set_min_num_points(static_cast<float>(parameter_dictionary->GetNonNegativeInt("min_num_points")));
  options.set_max_range(static_cast<float>(parameter_dictionary->GetDouble("max_range")));
  return options;
}

// cc:264-269
PointCloud AdaptiveVoxelFilter(const PointCloud& point_cloud, const proto::AdaptiveVoxelFilterOptions& options) {
  return AdaptivelyVoxelFiltered(options, FilterByMaxRange(point_cloud, options.max_range()));
}

}  // namespace sensor
}  // namespace cartographer
