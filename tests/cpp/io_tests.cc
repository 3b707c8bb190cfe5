// TEST INFRASTRUCTURE.  Compiles against the REPLACEMENT headers under cartographer-1_b200/host and runs
// the bodies of the reference's own unit tests for the host-side pieces next to the path, with the
// reference tests' inputs and expectations:
//   io/proto_stream_test.cc:38-63                  WriteAndReadBack
//   io/proto_stream_deserializer_test.cc:62-140    WorksOnGoldenTextStream, FailsIfVersionNotSupported
//   sensor/compressed_point_cloud_test.cc:50-118   CompressesPointsCorrectly, Compresses,
//                                                  CompressesEmptyPointCloud, CompressesNoGaps
//   sensor/internal/voxel_filter_test.cc:30-83     ReturnsOnePointInEachVoxel, HandlesLargeCoordinates,
//                                                  IgnoresTime (positions only)
// libprotobuf is absent: the few messages the tests need are written with the wire format's own rules
// (tag = field << 3 | wire type, varints, length-delimited sub-messages).  No GPU needed.
// Build and run: tests/test_host_cpp_cpu.py.
#include <sys/wait.h>
#include <unistd.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <string>
#include <vector>

#include "cartographer/io/proto_stream.h"
#include "cartographer/io/proto_stream_deserializer.h"
#include "cartographer/io/proto_wire.h"
#include "cartographer/sensor/compressed_point_cloud.h"
#include "cartographer/sensor/internal/voxel_filter.h"

using namespace cartographer;

static int g_checks = 0, g_failures = 0;
#define EXPECT_TRUE(c)                                                              \
  do {                                                                              \
    ++g_checks;                                                                     \
    if (!(c)) {                                                                     \
      ++g_failures;                                                                 \
      if (g_failures < 20) std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); \
    }                                                                               \
  } while (0)
#define EXPECT_FALSE(c) EXPECT_TRUE(!(c))
#define EXPECT_EQ(a, b) EXPECT_TRUE((a) == (b))

static std::string g_dir = ".";

// ---- wire-format writing, enough for the test messages ----
static std::string Varint(uint64_t v) {
  std::string out;
  while (v >= 0x80) {
    out.push_back(static_cast<char>((v & 0x7F) | 0x80));
    v >>= 7;
  }
  out.push_back(static_cast<char>(v));
  return out;
}
static std::string VarintField(int field, uint64_t v) { return Varint(static_cast<uint64_t>(field) << 3) + Varint(v); }
static std::string MessageField(int field, const std::string& body) {
  return Varint((static_cast<uint64_t>(field) << 3) | 2) + Varint(body.size()) + body;
}

// io/proto_stream_test.cc:38-63.  mapping::proto::Trajectory {repeated Node node = 1}, Node {int64 timestamp = 1}.
static void TestWriteAndReadBack() {
  const std::string test_file = g_dir + "/test_trajectory.pbstream";
  {
    io::ProtoStreamWriter writer(test_file);
    for (int i = 0; i != 10; ++i) writer.Write(MessageField(1, VarintField(1, static_cast<uint64_t>(i))));
    EXPECT_TRUE(writer.Close());
  }
  {
    io::ProtoStreamReader reader(test_file);
    for (int i = 0; i != 10; ++i) {
      std::string trajectory;
      EXPECT_TRUE(reader.Read(&trajectory));
      io::wire::Reader r(trajectory.data(), trajectory.size());
      int field = 0, type = 0, node_size = 0;
      int64_t timestamp = -1;
      while (r.Next(&field, &type)) {
        if (field == 1 && type == 2) {
          ++node_size;
          io::wire::Reader node(r.Bytes());
          int f = 0, t = 0;
          while (node.Next(&f, &t)) {
            if (f == 1 && t == 0) timestamp = static_cast<int64_t>(node.Varint());
            else node.Skip(t);
          }
        } else {
          r.Skip(type);
        }
      }
      EXPECT_EQ(1, node_size);
      EXPECT_EQ(i, timestamp);
    }
    std::string trajectory;
    EXPECT_FALSE(reader.Read(&trajectory));
    EXPECT_TRUE(reader.eof());
  }
  std::remove(test_file.c_str());
}

// The golden text stream of proto_stream_deserializer_test.cc:40-54 in wire form: "format_version: N";
// "pose_graph {}" ... "landmark_data {}" = SerializedData with the empty member 1 ... 9.
static void WriteGoldenStream(const std::string& file, int format_version, const std::vector<int>& members) {
  io::ProtoStreamWriter writer(file);
  writer.Write(VarintField(1, static_cast<uint64_t>(format_version)));
  for (const int member : members) writer.Write(MessageField(member, ""));
  EXPECT_TRUE(writer.Close());
}

// io/proto_stream_deserializer_test.cc:62-133
static void TestWorksOnGoldenTextStream() {
  const std::string file = g_dir + "/golden.pbstream";
  WriteGoldenStream(file, 1, {1, 2, 3, 4, 5, 6, 7, 8, 9});
  io::ProtoStreamReader reader(file);
  mapping::ValueConversionTables tables;
  io::ProtoStreamDeserializer deserializer(&reader, &tables);
  EXPECT_EQ(1u, deserializer.format_version());
  EXPECT_TRUE(deserializer.pose_graph().constraint.empty());
  EXPECT_TRUE(deserializer.pose_graph().trajectory.empty());
  EXPECT_EQ(0, deserializer.options_with_sensor_ids_size());
  io::SerializedData serialized_data;
  const io::SerializedData::DataCase expected[] = {
      io::SerializedData::kSubmap,          io::SerializedData::kNode,
      io::SerializedData::kTrajectoryData,  io::SerializedData::kImuData,
      io::SerializedData::kOdometryData,    io::SerializedData::kFixedFramePoseData,
      io::SerializedData::kLandmarkData};
  for (const io::SerializedData::DataCase data_case : expected) {
    EXPECT_TRUE(deserializer.ReadNextSerializedData(&serialized_data));
    EXPECT_EQ(data_case, serialized_data.data_case());
  }
  // "submap {}" / "node {}": every field at its proto3 default
  EXPECT_FALSE(deserializer.ReadNextSerializedData(&serialized_data));
  EXPECT_TRUE(reader.eof());
  std::remove(file.c_str());
}

// Runs `fn` in a child; true when the child died on SIGABRT (the replacement's CHECKs, as EXPECT_DEATH).
template <typename F>
static bool DiesOnCheck(const F& fn) {
  const pid_t pid = fork();
  if (pid == 0) {
    std::fclose(stderr);
    fn();
    _exit(0);
  }
  int status = 0;
  waitpid(pid, &status, 0);
  return WIFSIGNALED(status) && WTERMSIG(status) == SIGABRT;
}

// io/proto_stream_deserializer_test.cc:135-140 (format_version: 123), plus the stream-order CHECKs of
// proto_stream_deserializer.cc:56-70.
static void TestFailsIfVersionNotSupportedOrOutOfOrder() {
  const std::string file = g_dir + "/bad.pbstream";
  auto load = [&] {
    io::ProtoStreamReader reader(file);
    mapping::ValueConversionTables tables;
    io::ProtoStreamDeserializer deserializer(&reader, &tables);
  };
  WriteGoldenStream(file, 123, {});
  EXPECT_TRUE(DiesOnCheck(load));
  WriteGoldenStream(file, 2, {2, 1});  // options before the pose graph
  EXPECT_TRUE(DiesOnCheck(load));
  WriteGoldenStream(file, 2, {1});  // options missing
  EXPECT_TRUE(DiesOnCheck(load));
  WriteGoldenStream(file, 2, {1, 2});
  EXPECT_FALSE(DiesOnCheck(load));
  std::remove(file.c_str());
}

// ---- sensor/compressed_point_cloud_test.cc ----
constexpr float kPrecision = 0.001f;

static bool ApproximatelyEquals(const sensor::RangefinderPoint& point, const sensor::Vector3f& expected) {
  // Eigen's (a - b).isZero(kPrecision): every |coefficient| <= kPrecision
  return std::abs(point.position.x() - expected.x()) <= kPrecision &&
         std::abs(point.position.y() - expected.y()) <= kPrecision &&
         std::abs(point.position.z() - expected.z()) <= kPrecision;
}
static sensor::PointCloud CloudOf(std::initializer_list<sensor::Vector3f> points) {
  sensor::PointCloud cloud;
  for (const sensor::Vector3f& p : points) cloud.push_back({p});
  return cloud;
}

// :50-58
static void TestPoint(const sensor::Vector3f& p) {
  const sensor::CompressedPointCloud compressed(CloudOf({p}));
  EXPECT_EQ(1u, compressed.size());
  const sensor::PointCloud first = compressed.Decompress();
  EXPECT_TRUE(ApproximatelyEquals(first[0], p));
  const sensor::CompressedPointCloud recompressed(CloudOf({first[0].position}));
  EXPECT_TRUE(ApproximatelyEquals(recompressed.Decompress()[0], p));
}

// :60-72
static void TestCompressesPointsCorrectly() {
  TestPoint(sensor::Vector3f(8000.f, 7500.f, 5000.f));
  TestPoint(sensor::Vector3f(1000.f, 2000.f, 3000.f));
  TestPoint(sensor::Vector3f(100.f, 200.f, 300.f));
  TestPoint(sensor::Vector3f(10.f, 20.f, 30.f));
  TestPoint(sensor::Vector3f(-0.00049f, -0.0005f, -0.0015f));
  TestPoint(sensor::Vector3f(0.05119f, 0.0512f, 0.05121f));
  TestPoint(sensor::Vector3f(-0.05119f, -0.0512f, -0.05121f));
  TestPoint(sensor::Vector3f(0.8405f, 0.84f, 0.8396f));
  TestPoint(sensor::Vector3f(0.8395f, 0.8394f, 0.8393f));
  TestPoint(sensor::Vector3f(0.839f, 0.8391f, 0.8392f));
  TestPoint(sensor::Vector3f(0.8389f, 0.8388f, 0.83985f));
}

// :74-89
static void TestCompresses() {
  const sensor::CompressedPointCloud compressed(
      CloudOf({sensor::Vector3f(0.838f, 0, 0), sensor::Vector3f(0.839f, 0, 0), sensor::Vector3f(0.840f, 0, 0)}));
  EXPECT_FALSE(compressed.empty());
  EXPECT_EQ(3u, compressed.size());
  const sensor::PointCloud decompressed = compressed.Decompress();
  EXPECT_EQ(3u, decompressed.size());
  for (const float x : {0.838f, 0.839f, 0.840f}) {
    bool contains = false;
    for (const sensor::RangefinderPoint& p : decompressed) contains |= ApproximatelyEquals(p, sensor::Vector3f(x, 0, 0));
    EXPECT_TRUE(contains);
  }
}

// :91-95
static void TestCompressesEmptyPointCloud() {
  const sensor::CompressedPointCloud compressed;
  EXPECT_TRUE(compressed.empty());
  EXPECT_EQ(0u, compressed.size());
}

// :101-118
static void TestCompressesNoGaps() {
  sensor::PointCloud point_cloud;
  for (int i = 0; i < 3000; ++i) point_cloud.push_back({sensor::Vector3f(kPrecision * i - 1.5f, 0, 0)});
  const sensor::CompressedPointCloud compressed(point_cloud);
  const sensor::PointCloud decompressed = compressed.Decompress();
  const sensor::CompressedPointCloud recompressed(decompressed);
  EXPECT_EQ(decompressed.size(), recompressed.size());
  std::vector<float> x_coord;
  for (const sensor::RangefinderPoint& p : decompressed) x_coord.push_back(p.position.x());
  std::sort(x_coord.begin(), x_coord.end());
  for (size_t i = 1; i < x_coord.size(); ++i)
    EXPECT_TRUE(std::abs(std::abs(x_coord[i] - x_coord[i - 1]) - kPrecision) <= 1e-7f);
}

// ---- sensor/internal/voxel_filter_test.cc ----
static bool Contains(const sensor::PointCloud& cloud, const sensor::RangefinderPoint& p) {
  for (const sensor::RangefinderPoint& q : cloud)
    if (q.position.x() == p.position.x() && q.position.y() == p.position.y() && q.position.z() == p.position.z())
      return true;
  return false;
}

// :30-42
static void TestReturnsOnePointInEachVoxel() {
  const sensor::PointCloud point_cloud = CloudOf({sensor::Vector3f(0.f, 0.f, 0.f), sensor::Vector3f(0.1f, -0.1f, 0.1f),
                                                  sensor::Vector3f(0.3f, -0.1f, 0.f), sensor::Vector3f(0.f, 0.f, 0.1f)});
  const sensor::PointCloud result = sensor::VoxelFilter(point_cloud, 0.3f);
  EXPECT_EQ(result.size(), 2u);
  EXPECT_TRUE(Contains(point_cloud, result[0]));
  EXPECT_TRUE(Contains(point_cloud, result[1]));
  EXPECT_TRUE(Contains(result, point_cloud[2]));
}

// :63-73
static void TestHandlesLargeCoordinates() {
  const sensor::PointCloud point_cloud =
      CloudOf({sensor::Vector3f(100000.f, 0.f, 0.f), sensor::Vector3f(100000.001f, -0.0001f, 0.0001f),
               sensor::Vector3f(100000.003f, -0.0001f, 0.f), sensor::Vector3f(-200000.f, 0.f, 0.f)});
  const sensor::PointCloud result = sensor::VoxelFilter(point_cloud, 0.01f);
  EXPECT_EQ(result.size(), 2u);
  EXPECT_TRUE(Contains(result, point_cloud[3]));
}

// :75-83 (the timed overload's property on the untimed one: 100 points of one voxel leave one)
static void TestManyPointsInOneVoxel() {
  std::vector<sensor::RangefinderPoint> points;
  for (int i = 0; i < 100; ++i) points.push_back({sensor::Vector3f(-100.f, 0.3f, 0.4f)});
  const std::vector<sensor::RangefinderPoint> result = sensor::VoxelFilter(points, 0.3f);
  EXPECT_EQ(result.size(), 1u);
}

// CreateAdaptiveVoxelFilterOptions reads the three keys of the Lua block (voxel_filter.cc:253-262).
static void TestAdaptiveOptions() {
  common::LuaParameterDictionary dictionary("return { max_length = 0.9, min_num_points = 100, max_range = 50. }");
  const sensor::proto::AdaptiveVoxelFilterOptions options = sensor::CreateAdaptiveVoxelFilterOptions(&dictionary);
  EXPECT_EQ(options.max_length(), 0.9f);
  EXPECT_EQ(options.min_num_points(), 100.f);
  EXPECT_EQ(options.max_range(), 50.f);
  sensor::PointCloud cloud;
  for (int i = 0; i < 1000; ++i) cloud.push_back({sensor::Vector3f(0.01f * i, 0.f, 0.f)});
  cloud.push_back({sensor::Vector3f(60.f, 0.f, 0.f)});  // beyond max_range
  const sensor::PointCloud filtered = sensor::AdaptiveVoxelFilter(cloud, options);
  EXPECT_TRUE(filtered.size() >= 100 && filtered.size() < 1000);
  for (const sensor::RangefinderPoint& p : filtered) EXPECT_TRUE(p.position.x() <= 50.f);
}

int main(int argc, char** argv) {
  if (argc > 1) g_dir = argv[1];
  TestWriteAndReadBack();
  TestWorksOnGoldenTextStream();
  TestFailsIfVersionNotSupportedOrOutOfOrder();
  TestCompressesPointsCorrectly();
  TestCompresses();
  TestCompressesEmptyPointCloud();
  TestCompressesNoGaps();
  TestReturnsOnePointInEachVoxel();
  TestHandlesLargeCoordinates();
  TestManyPointsInOneVoxel();
  TestAdaptiveOptions();
  std::printf("%d checks, %d failures\n", g_checks, g_failures);
  return g_failures == 0 ? 0 : 1;
}
