// Flat C entry points over the host-side point-cloud filters (cartographer/sensor/internal/
// voxel_filter.h), for ctypes.  Product code: never touches oracle/.
#include <cstdint>
#include <vector>

#include "cartographer/sensor/compressed_point_cloud.h"
#include "cartographer/sensor/internal/voxel_filter.h"

using namespace cartographer;

namespace {
sensor::PointCloud MakeCloud(const float* xyz, int n) {
  std::vector<sensor::RangefinderPoint> points(n);
  for (int i = 0; i < n; ++i) points[i].position = sensor::Vector3f(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
  return sensor::PointCloud(std::move(points));
}
int StoreCloud(const sensor::PointCloud& cloud, float* out_xyz) {
  for (size_t i = 0; i < cloud.size(); ++i) {
    out_xyz[3 * i] = cloud[i].position.x();
    out_xyz[3 * i + 1] = cloud[i].position.y();
    out_xyz[3 * i + 2] = cloud[i].position.z();
  }
  return static_cast<int>(cloud.size());
}
}  // namespace

extern "C" {

// which: 0 VoxelFilter(cloud, a); 1 AdaptiveVoxelFilter(cloud, {max_length a, min_num_points b,
// max_range c}); 2 VoxelFilterEdge(cloud, rad a, voxeledgeratio b).  out_xyz holds up to n points;
// returns the number kept.
int csmh_voxel_filter(int which, const float* xyz, int n, float a, float b, float c, float* out_xyz) {
  const sensor::PointCloud cloud = MakeCloud(xyz, n);
  if (which == 0) return StoreCloud(sensor::VoxelFilter(cloud, a), out_xyz);
  if (which == 1) {
    sensor::proto::AdaptiveVoxelFilterOptions options;
    options.set_max_length(a);
    options.set_min_num_points(b);
    options.set_max_range(c);
    return StoreCloud(sensor::AdaptiveVoxelFilter(cloud, options), out_xyz);
  }
  return StoreCloud(sensor::VoxelFilterEdge(cloud, a, b), out_xyz);
}

// sensor::CompressedPointCloud(cloud).point_data(): returns the number of int32 words (out may be
// NULL to count; -1 when `cap` is too small).
int64_t csmh_compress_point_cloud(const float* xyz, int n, int32_t* out, int64_t cap) {
  const sensor::CompressedPointCloud compressed(MakeCloud(xyz, n));
  const std::vector<int32_t>& words = compressed.point_data();
  if (out != nullptr) {
    if (static_cast<int64_t>(words.size()) > cap) return -1;
    for (size_t i = 0; i < words.size(); ++i) out[i] = words[i];
  }
  return static_cast<int64_t>(words.size());
}

}  // extern "C"
