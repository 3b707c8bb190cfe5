// TEST INFRASTRUCTURE ONLY (never linked or loaded by the product).  C surface over the
// reference's OWN .pbstream container and point-cloud compression, compiled unmodified from
// /root/reference (oracle/Makefile `ref`):
//   io::ProtoStreamWriter / ProtoStreamReader           cartographer/io/proto_stream.cc
//   common::FastGzipString / FastGunzipString           cartographer/common/port.h:48-66
//   sensor::CompressedPointCloud                        cartographer/sensor/compressed_point_cloud.cc
// libprotobuf is absent, so messages enter and leave as already-serialized bytes inside a
// stand-in google::protobuf::Message (ref_shim/google/protobuf/message.h); boost's gzip filters
// are a zlib-backed stand-in (ref_shim/boost/iostreams/filtering_stream.hpp).  What the tests pin
// with this: the container bytes the product's reader must accept and its writer must produce,
// and the decoded floats of a compressed point cloud.
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "cartographer/io/proto_stream.h"
#include "cartographer/sensor/compressed_point_cloud.h"

namespace {

class RawMessage : public google::protobuf::Message {
 public:
  bool SerializeToString(std::string* output) const override {
    *output = bytes;
    return true;
  }
  bool ParseFromString(const std::string& data) override {
    bytes = data;
    return true;
  }
  std::string bytes;
};

}  // namespace

extern "C" {

// `count` serialized messages (message i = bytes[offsets[i] .. offsets[i + 1])) -> a .pbstream
// written by the reference's ProtoStreamWriter.
int oracle_pbstream_write(const char* filename, int count, const uint8_t* bytes, const int64_t* offsets) {
  cartographer::io::ProtoStreamWriter writer(filename);
  for (int i = 0; i < count; ++i) {
    RawMessage m;
    m.bytes.assign(reinterpret_cast<const char*>(bytes) + offsets[i], static_cast<size_t>(offsets[i + 1] - offsets[i]));
    writer.WriteProto(m);
  }
  return writer.Close() ? 0 : -1;
}

// Every message of a .pbstream through the reference's ProtoStreamReader, concatenated into
// `bytes` (capacity cap) with offsets[count + 1]; returns the count, or -(bytes needed) - 1.
int64_t oracle_pbstream_read(const char* filename, uint8_t* bytes, int64_t cap, int64_t* offsets, int max_count) {
  cartographer::io::ProtoStreamReader reader(filename);
  std::vector<std::string> messages;
  int64_t total = 0;
  for (;;) {
    RawMessage m;
    if (!reader.ReadProto(&m)) break;
    total += static_cast<int64_t>(m.bytes.size());
    messages.push_back(m.bytes);
  }
  if (total > cap || static_cast<int64_t>(messages.size()) > max_count) return -total - 1;
  int64_t at = 0;
  for (size_t i = 0; i < messages.size(); ++i) {
    offsets[i] = at;
    std::memcpy(bytes + at, messages[i].data(), messages[i].size());
    at += static_cast<int64_t>(messages[i].size());
  }
  offsets[messages.size()] = at;
  return static_cast<int64_t>(messages.size());
}

// common::FastGzipString / FastGunzipString on a buffer; returns the output size (or -size - 1
// when cap is too small).
int64_t oracle_gzip(const uint8_t* in, int64_t n, int decompress, uint8_t* out, int64_t cap) {
  const std::string input(reinterpret_cast<const char*>(in), static_cast<size_t>(n));
  std::string output;
  if (decompress) cartographer::common::FastGunzipString(input, &output);
  else cartographer::common::FastGzipString(input, &output);
  if (static_cast<int64_t>(output.size()) > cap) return -static_cast<int64_t>(output.size()) - 1;
  std::memcpy(out, output.data(), output.size());
  return static_cast<int64_t>(output.size());
}

// CompressedPointCloud(PointCloud).ToProto().point_data(): returns the number of int32 words
// (out may be NULL to count).
int64_t oracle_compress_point_cloud(const float* xyz, int n, int32_t* out, int64_t cap) {
  std::vector<cartographer::sensor::RangefinderPoint> points(n);
  for (int i = 0; i < n; ++i) points[i].position = Eigen::Vector3f(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
  const cartographer::sensor::CompressedPointCloud compressed{cartographer::sensor::PointCloud(points)};
  const cartographer::sensor::proto::CompressedPointCloud proto = compressed.ToProto();
  if (out != nullptr) {
    if (proto.point_data_size() > cap) return -1;
    for (int i = 0; i < proto.point_data_size(); ++i) out[i] = proto.point_data(i);
  }
  return proto.point_data_size();
}

// CompressedPointCloud(proto).Decompress() -> xyz[num_points * 3].
void oracle_decompress_point_cloud(const int32_t* point_data, int64_t words, int num_points, float* xyz) {
  cartographer::sensor::proto::CompressedPointCloud proto;
  proto.set_num_points(num_points);
  for (int64_t i = 0; i < words; ++i) proto.add_point_data(point_data[i]);
  const cartographer::sensor::PointCloud cloud = cartographer::sensor::CompressedPointCloud(proto).Decompress();
  for (size_t i = 0; i < cloud.size(); ++i) {
    xyz[3 * i] = cloud[i].position.x();
    xyz[3 * i + 1] = cloud[i].position.y();
    xyz[3 * i + 2] = cloud[i].position.z();
  }
}

}  // extern "C"

// ---- the reference's own voxel filters (cartographer/sensor/internal/voxel_filter.cc) ----
#include "cartographer/sensor/internal/voxel_filter.h"

namespace {
cartographer::sensor::PointCloud MakeCloud(const float* xyz, int n) {
  std::vector<cartographer::sensor::RangefinderPoint> points(n);
  for (int i = 0; i < n; ++i) points[i].position = Eigen::Vector3f(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
  return cartographer::sensor::PointCloud(points);
}
int StoreCloud(const cartographer::sensor::PointCloud& cloud, float* out_xyz) {
  for (size_t i = 0; i < cloud.size(); ++i) {
    out_xyz[3 * i] = cloud[i].position.x();
    out_xyz[3 * i + 1] = cloud[i].position.y();
    out_xyz[3 * i + 2] = cloud[i].position.z();
  }
  return static_cast<int>(cloud.size());
}
}  // namespace

extern "C" {
// which: 0 VoxelFilter(cloud, a), 1 AdaptiveVoxelFilter(cloud, {max_length a, min_num_points b,
// max_range c}), 2 VoxelFilterEdge(cloud, rad a, voxeledgeratio b).  out_xyz holds up to n points;
// returns the number kept.
int oracle_voxel_filter(int which, const float* xyz, int n, float a, float b, float c, float* out_xyz) {
  const cartographer::sensor::PointCloud cloud = MakeCloud(xyz, n);
  if (which == 0) return StoreCloud(cartographer::sensor::VoxelFilter(cloud, a), out_xyz);
  if (which == 1) {
    cartographer::sensor::proto::AdaptiveVoxelFilterOptions options;
    options.set_max_length(a);
    options.set_min_num_points(b);
    options.set_max_range(c);
    return StoreCloud(cartographer::sensor::AdaptiveVoxelFilter(cloud, options), out_xyz);
  }
  return StoreCloud(cartographer::sensor::VoxelFilterEdge(cloud, a, b), out_xyz);
}
}  // extern "C"
