// Minimal stand-ins for the Cartographer value types the scan-matching boundary
// uses.  This image has no Eigen / protobuf / glog, so the host layer compiles
// against these; in a full Cartographer tree the real headers replace this file
// (same names, same accessors, same semantics):
//   transform::Rigid2d            cartographer/transform/rigid_transform.h:34-103
//   sensor::PointCloud            cartographer/sensor/point_cloud.h:33-92
//   mapping::CellLimits           cartographer/mapping/2d/xy_index.h:33-45
//   mapping::MapLimits            cartographer/mapping/2d/map_limits.h:39-98
//   mapping::ValueConversionTables cartographer/mapping/value_conversion_tables.{h,cc}
//   mapping::Grid2D/ProbabilityGrid cartographer/mapping/2d/grid_2d.h:38-128,
//                                  cartographer/mapping/2d/probability_grid.{h,cc}
//   proto::*Options               cartographer/mapping/proto/scan_matching/*.proto,
//                                  cartographer/mapping/proto/pose_graph/constraint_builder_options.proto
#ifndef CSM_HOST_CARTOGRAPHER_COMPAT_H_
#define CSM_HOST_CARTOGRAPHER_COMPAT_H_

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <memory>
#include <string>
#include <tuple>
#include <vector>

#define CSM_CHECK(cond)                                                               \
  do {                                                                                \
    if (!(cond)) {                                                                    \
      std::fprintf(stderr, "Check failed: %s (%s:%d)\n", #cond, __FILE__, __LINE__);  \
      std::abort();                                                                   \
    }                                                                                 \
  } while (0)

// The subset of Eigen the public helpers of the replaced headers name in their signatures
// (Array2i cell indices, Translation2f).  In a full Cartographer tree the real Eigen provides them.
namespace Eigen {
struct Array2b {
  bool v[2];
  bool all() const { return v[0] && v[1]; }
  bool any() const { return v[0] || v[1]; }
};
struct Array2i {
  int v[2] = {0, 0};
  Array2i() = default;
  Array2i(int x, int y) : v{x, y} {}
  static Array2i Zero() { return Array2i(0, 0); }
  int x() const { return v[0]; }
  int y() const { return v[1]; }
  int& x() { return v[0]; }
  int& y() { return v[1]; }
  Array2i operator+(const Array2i& o) const { return Array2i(v[0] + o.v[0], v[1] + o.v[1]); }
  Array2i operator-(const Array2i& o) const { return Array2i(v[0] - o.v[0], v[1] - o.v[1]); }
  Array2b operator==(const Array2i& o) const { return Array2b{{v[0] == o.v[0], v[1] == o.v[1]}}; }
};
struct Translation2f {
  float v[2] = {0.f, 0.f};
  Translation2f() = default;
  Translation2f(float x, float y) : v{x, y} {}
  static Translation2f Identity() { return Translation2f(); }
  float x() const { return v[0]; }
  float y() const { return v[1]; }
};
}  // namespace Eigen

namespace cartographer {

using uint8 = uint8_t;
using uint16 = uint16_t;

namespace common {
inline int RoundToInt(const float x) { return static_cast<int>(std::lround(x)); }
inline int RoundToInt(const double x) { return static_cast<int>(std::lround(x)); }
}  // namespace common

namespace transform {
struct Vector2d {
  double v[2] = {0., 0.};
  Vector2d() = default;
  Vector2d(double x, double y) : v{x, y} {}
  double x() const { return v[0]; }
  double y() const { return v[1]; }
  double norm() const { return std::sqrt(v[0] * v[0] + v[1] * v[1]); }
};
// Rotation kept as its angle, like Eigen::Rotation2Dd.
class Rotation2Dd {
 public:
  Rotation2Dd() = default;
  explicit Rotation2Dd(double angle) : angle_(angle) {}
  double angle() const { return angle_; }
  Rotation2Dd operator*(const Rotation2Dd& o) const { return Rotation2Dd(angle_ + o.angle_); }
  Rotation2Dd inverse() const { return Rotation2Dd(-angle_); }
  Vector2d operator*(const Vector2d& p) const {
    const double s = std::sin(angle_), c = std::cos(angle_);
    return Vector2d(c * p.x() + -s * p.y(), s * p.x() + c * p.y());
  }

 private:
  double angle_ = 0.;
};
class Rigid2d {
 public:
  Rigid2d() = default;
  Rigid2d(const Vector2d& t, const Rotation2Dd& r) : translation_(t), rotation_(r) {}
  Rigid2d(const Vector2d& t, double r) : translation_(t), rotation_(r) {}
  static Rigid2d Identity() { return Rigid2d(); }
  static Rigid2d Translation(const Vector2d& t) { return Rigid2d(t, Rotation2Dd()); }
  static Rigid2d Rotation(double r) { return Rigid2d(Vector2d(), r); }
  const Vector2d& translation() const { return translation_; }
  Rotation2Dd rotation() const { return rotation_; }
  Rigid2d inverse() const {
    const Rotation2Dd r = rotation_.inverse();
    const Vector2d t = r * translation_;
    return Rigid2d(Vector2d(-t.x(), -t.y()), r);
  }

 private:
  Vector2d translation_;
  Rotation2Dd rotation_;
};
inline Rigid2d operator*(const Rigid2d& l, const Rigid2d& r) {
  const Vector2d t = l.rotation() * r.translation();
  return Rigid2d(Vector2d(t.x() + l.translation().x(), t.y() + l.translation().y()),
                 l.rotation() * r.rotation());
}

// Minimal Rigid3d (translation + unit quaternion w, x, y, z), enough for
// Embed3D / Project2D (reference: transform/transform.h:34-115,
// transform/rigid_transform.h:116-196).
struct Vector3d {
  double v[3] = {0., 0., 0.};
  Vector3d() = default;
  Vector3d(double x, double y, double z) : v{x, y, z} {}
  double x() const { return v[0]; }
  double y() const { return v[1]; }
  double z() const { return v[2]; }
};
struct Quaterniond {
  double w_ = 1., x_ = 0., y_ = 0., z_ = 0.;
  Quaterniond() = default;
  Quaterniond(double w, double x, double y, double z) : w_(w), x_(x), y_(y), z_(z) {}
  double w() const { return w_; }
  double x() const { return x_; }
  double y() const { return y_; }
  double z() const { return z_; }
};
class Rigid3d {
 public:
  Rigid3d() = default;
  Rigid3d(const Vector3d& t, const Quaterniond& r) : translation_(t), rotation_(r) {}
  static Rigid3d Identity() { return Rigid3d(); }
  const Vector3d& translation() const { return translation_; }
  const Quaterniond& rotation() const { return rotation_; }

 private:
  Vector3d translation_;
  Quaterniond rotation_;
};
// transform.h:86-91: AngleAxis(rotation.angle(), UnitZ) -> quaternion.
inline Rigid3d Embed3D(const Rigid2d& t) {
  const double half = 0.5 * t.rotation().angle();
  return Rigid3d(Vector3d(t.translation().x(), t.translation().y(), 0.),
                 Quaterniond(std::cos(half), 0., 0., std::sin(half)));
}
// transform.h:57-61: yaw of the rotated x axis.
inline double GetYaw(const Quaterniond& q) {
  // direction = q * UnitX
  const double dx = 1. - 2. * (q.y() * q.y() + q.z() * q.z());
  const double dy = 2. * (q.x() * q.y() + q.w() * q.z());
  return std::atan2(dy, dx);
}
// transform.h:79-84
inline Rigid2d Project2D(const Rigid3d& t) {
  return Rigid2d(Vector2d(t.translation().x(), t.translation().y()), GetYaw(t.rotation()));
}
}  // namespace transform

namespace sensor {
struct Vector3f {
  float v[3] = {0.f, 0.f, 0.f};
  Vector3f() = default;
  Vector3f(float x, float y, float z) : v{x, y, z} {}
  static Vector3f Zero() { return Vector3f(); }
  float x() const { return v[0]; }
  float y() const { return v[1]; }
  float z() const { return v[2]; }
};
struct RangefinderPoint {
  Vector3f position;
};
class PointCloud {
 public:
  PointCloud() = default;
  explicit PointCloud(std::vector<RangefinderPoint> p) : points_(std::move(p)) {}
  size_t size() const { return points_.size(); }
  bool empty() const { return points_.empty(); }
  const std::vector<RangefinderPoint>& points() const { return points_; }
  const RangefinderPoint& operator[](size_t i) const { return points_[i]; }
  void push_back(RangefinderPoint p) { points_.push_back(p); }
  std::vector<RangefinderPoint>::const_iterator begin() const { return points_.begin(); }
  std::vector<RangefinderPoint>::const_iterator end() const { return points_.end(); }

 private:
  std::vector<RangefinderPoint> points_;
};
// sensor/range_data.h:32-36
struct RangeData {
  Vector3f origin;
  PointCloud returns;
  PointCloud misses;
};
}  // namespace sensor

namespace mapping {

struct CellLimits {
  CellLimits() = default;
  CellLimits(int x, int y) : num_x_cells(x), num_y_cells(y) {}
  int num_x_cells = 0;
  int num_y_cells = 0;
};

// mapping/2d/xy_index.h:49-110: row-major walk over a range of cell indices (x fastest).
class XYIndexRangeIterator {
 public:
  XYIndexRangeIterator(const Eigen::Array2i& min_xy_index, const Eigen::Array2i& max_xy_index)
      : min_xy_index_(min_xy_index), max_xy_index_(max_xy_index), xy_index_(min_xy_index) {}
  explicit XYIndexRangeIterator(const CellLimits& cell_limits)
      : XYIndexRangeIterator(Eigen::Array2i::Zero(),
                             Eigen::Array2i(cell_limits.num_x_cells - 1, cell_limits.num_y_cells - 1)) {}
  XYIndexRangeIterator& operator++() {
    if (xy_index_.x() < max_xy_index_.x()) {
      ++xy_index_.x();
    } else {
      xy_index_.x() = min_xy_index_.x();
      ++xy_index_.y();
    }
    return *this;
  }
  Eigen::Array2i& operator*() { return xy_index_; }
  bool operator==(const XYIndexRangeIterator& other) const { return (xy_index_ == other.xy_index_).all(); }
  bool operator!=(const XYIndexRangeIterator& other) const { return !operator==(other); }
  XYIndexRangeIterator begin() { return XYIndexRangeIterator(min_xy_index_, max_xy_index_); }
  XYIndexRangeIterator end() {
    XYIndexRangeIterator it = begin();
    it.xy_index_ = Eigen::Array2i(min_xy_index_.x(), max_xy_index_.y() + 1);
    return it;
  }

 private:
  Eigen::Array2i min_xy_index_, max_xy_index_, xy_index_;
};

class MapLimits {
 public:
  MapLimits(double resolution, const transform::Vector2d& max, const CellLimits& cell_limits)
      : resolution_(resolution), max_(max), cell_limits_(cell_limits) {
    CSM_CHECK(resolution_ > 0.);
    CSM_CHECK(cell_limits.num_x_cells > 0);
    CSM_CHECK(cell_limits.num_y_cells > 0);
  }
  double resolution() const { return resolution_; }
  const transform::Vector2d& max() const { return max_; }
  const CellLimits& cell_limits() const { return cell_limits_; }
  // map_limits.h:84-90
  bool Contains(const Eigen::Array2i& cell_index) const {
    return cell_index.x() >= 0 && cell_index.y() >= 0 && cell_index.x() < cell_limits_.num_x_cells &&
           cell_index.y() < cell_limits_.num_y_cells;
  }

 private:
  double resolution_;
  transform::Vector2d max_;
  CellLimits cell_limits_;
};

constexpr float kMinProbability = 0.1f;
constexpr float kMaxProbability = 1.f - kMinProbability;
constexpr float kMinCorrespondenceCost = 1.f - kMaxProbability;
constexpr float kMaxCorrespondenceCost = 1.f - kMinProbability;

// probability_values.h:32-46, :85-88, :32-41 of the .cc (fp32, as the reference)
inline uint16 BoundedFloatToValue(const float value, const float lower_bound, const float upper_bound) {
  const float clamped = std::min(std::max(value, lower_bound), upper_bound);
  return static_cast<uint16>(std::lround((clamped - lower_bound) * (32766.f / (upper_bound - lower_bound))) + 1);
}
inline uint16 CorrespondenceCostToValue(const float correspondence_cost) {
  return BoundedFloatToValue(correspondence_cost, kMinCorrespondenceCost, kMaxCorrespondenceCost);
}
inline float Odds(float probability) { return probability / (1.f - probability); }
inline float ProbabilityFromOdds(const float odds) { return odds / (odds + 1.f); }

// value_conversion_tables.cc:29-67
class ValueConversionTables {
 public:
  const std::vector<float>* GetConversionTable(float unknown_result, float lower_bound,
                                               float upper_bound) {
    const auto key = std::make_tuple(unknown_result, lower_bound, upper_bound);
    auto it = tables_.find(key);
    if (it != tables_.end()) return it->second.get();
    auto table = std::make_unique<std::vector<float>>();
    table->reserve(65536);
    const float scale = (upper_bound - lower_bound) / 32766.f;
    for (uint32_t v = 0; v != 65536; ++v) {
      const uint16 masked = static_cast<uint16>(v) & 0x7FFF;
      table->push_back(masked == 0 ? unknown_result : masked * scale + (lower_bound - scale));
    }
    return tables_.emplace(key, std::move(table)).first->second.get();
  }

 private:
  std::map<std::tuple<float, float, float>, std::unique_ptr<const std::vector<float>>> tables_;
};

enum class GridType { PROBABILITY_GRID, TSDF };

class Grid2D {
 public:
  Grid2D(const MapLimits& limits, float min_correspondence_cost, float max_correspondence_cost,
         ValueConversionTables* conversion_tables)
      : limits_(limits),
        correspondence_cost_cells_(
            static_cast<size_t>(limits.cell_limits().num_x_cells) * limits.cell_limits().num_y_cells, 0),
        min_correspondence_cost_(min_correspondence_cost),
        max_correspondence_cost_(max_correspondence_cost),
        value_to_correspondence_cost_table_(conversion_tables->GetConversionTable(
            max_correspondence_cost, min_correspondence_cost, max_correspondence_cost)) {
    CSM_CHECK(min_correspondence_cost_ < max_correspondence_cost_);
  }
  virtual ~Grid2D() {}
  const MapLimits& limits() const { return limits_; }
  float GetMinCorrespondenceCost() const { return min_correspondence_cost_; }
  float GetMaxCorrespondenceCost() const { return max_correspondence_cost_; }
  virtual GridType GetGridType() const = 0;
  // Accessors the device path needs in place of per-cell GetCorrespondenceCost().
  const std::vector<uint16>& correspondence_cost_cells() const { return correspondence_cost_cells_; }
  std::vector<uint16>* mutable_correspondence_cost_cells() { return &correspondence_cost_cells_; }
  const std::vector<float>& value_to_correspondence_cost_table() const {
    return *value_to_correspondence_cost_table_;
  }
  // grid_2d.cc:131-166: doubles the grid around its centre until it contains `point`.
  virtual void GrowLimits(const float px, const float py) {
    auto contains = [this](const float x, const float y) {
      const long cx = std::lround((limits_.max().y() - y) / limits_.resolution() - 0.5);
      const long cy = std::lround((limits_.max().x() - x) / limits_.resolution() - 0.5);
      return cx >= 0 && cy >= 0 && cx < limits_.cell_limits().num_x_cells &&
             cy < limits_.cell_limits().num_y_cells;
    };
    while (!contains(px, py)) {
      const int nx = limits_.cell_limits().num_x_cells, ny = limits_.cell_limits().num_y_cells;
      const int x_offset = nx / 2, y_offset = ny / 2;
      const MapLimits grown(limits_.resolution(),
                            transform::Vector2d(limits_.max().x() + limits_.resolution() * y_offset,
                                                limits_.max().y() + limits_.resolution() * x_offset),
                            CellLimits(2 * nx, 2 * ny));
      const int stride = 2 * nx;
      std::vector<uint16> cells(static_cast<size_t>(4) * nx * ny, 0);
      for (int i = 0; i < ny; ++i)
        for (int j = 0; j < nx; ++j)
          cells[x_offset + stride * y_offset + j + i * stride] = correspondence_cost_cells_[j + i * nx];
      correspondence_cost_cells_.swap(cells);
      limits_ = grown;
      Grown(nx, ny, x_offset, y_offset);
    }
  }

 protected:
  // Called after every doubling with the old size and the offset of the old cells in the new grid,
  // for grids that carry a second cell array (TSDF2D::GrowLimits, tsdf_2d.cc:100-105).
  virtual void Grown(int /*old_nx*/, int /*old_ny*/, int /*x_offset*/, int /*y_offset*/) {}

 private:
  MapLimits limits_;
  std::vector<uint16> correspondence_cost_cells_;
  float min_correspondence_cost_;
  float max_correspondence_cost_;
  const std::vector<float>* value_to_correspondence_cost_table_;
};

class ProbabilityGrid : public Grid2D {
 public:
  ProbabilityGrid(const MapLimits& limits, ValueConversionTables* conversion_tables)
      : Grid2D(limits, kMinCorrespondenceCost, kMaxCorrespondenceCost, conversion_tables) {}
  // What ProbabilityGrid(const proto::Grid2D&, ...) builds (probability_grid.cc:33-37): the
  // correspondence-cost bounds come from the message (io/grid_proto_reader.h).
  ProbabilityGrid(const MapLimits& limits, float min_correspondence_cost, float max_correspondence_cost,
                  ValueConversionTables* conversion_tables)
      : Grid2D(limits, min_correspondence_cost, max_correspondence_cost, conversion_tables) {}
  GridType GetGridType() const override { return GridType::PROBABILITY_GRID; }
  // probability_grid.cc:40-48: sets the probability of a cell that was unknown so far.
  void SetProbability(const Eigen::Array2i& cell_index, const float probability) {
    CSM_CHECK(limits().Contains(cell_index));
    uint16& cell = (*mutable_correspondence_cost_cells())[
        static_cast<size_t>(limits().cell_limits().num_x_cells) * cell_index.y() + cell_index.x()];
    CSM_CHECK(cell == 0);
    cell = CorrespondenceCostToValue(1.f - probability);
  }
  // probability_grid.cc:78-82
  float GetProbability(const Eigen::Array2i& cell_index) const {
    if (!limits().Contains(cell_index)) return kMinProbability;
    return 1.f - value_to_correspondence_cost_table()[correspondence_cost_cells()[
        static_cast<size_t>(limits().cell_limits().num_x_cells) * cell_index.y() + cell_index.x()]];
  }
};

// mapping/internal/2d/tsd_value_converter.{h,cc}
class TSDValueConverter {
 public:
  TSDValueConverter(float max_tsd, float max_weight, ValueConversionTables* conversion_tables)
      : max_tsd_(max_tsd),
        min_tsd_(-max_tsd),
        max_weight_(max_weight),
        tsd_resolution_(32766.f / (max_tsd_ - min_tsd_)),
        weight_resolution_(32766.f / (max_weight_ - min_weight_)),
        value_to_tsd_(conversion_tables->GetConversionTable(min_tsd_, min_tsd_, max_tsd_)),
        value_to_weight_(conversion_tables->GetConversionTable(min_weight_, min_weight_, max_weight)) {}
  uint16 TSDToValue(const float tsd) const {
    return static_cast<uint16>(
        std::lround((std::min(std::max(tsd, min_tsd_), max_tsd_) - min_tsd_) * tsd_resolution_) + 1);
  }
  uint16 WeightToValue(const float weight) const {
    return static_cast<uint16>(
        std::lround((std::min(std::max(weight, min_weight_), max_weight_) - min_weight_) * weight_resolution_) + 1);
  }
  float ValueToTSD(const uint16 value) const { return (*value_to_tsd_)[value]; }
  float ValueToWeight(const uint16 value) const { return (*value_to_weight_)[value]; }
  static uint16 getUnknownTSDValue() { return 0; }
  static uint16 getUnknownWeightValue() { return 0; }
  static uint16 getUpdateMarker() { return 1u << 15; }
  float getMaxTSD() const { return max_tsd_; }
  float getMinTSD() const { return min_tsd_; }
  float getMaxWeight() const { return max_weight_; }
  float getMinWeight() const { return min_weight_; }
  const std::vector<float>& value_to_tsd_table() const { return *value_to_tsd_; }
  const std::vector<float>& value_to_weight_table() const { return *value_to_weight_; }

 private:
  float max_tsd_, min_tsd_, max_weight_, tsd_resolution_, weight_resolution_;
  static constexpr float min_weight_ = 0.f;
  const std::vector<float>* value_to_tsd_;
  const std::vector<float>* value_to_weight_;
};

// mapping/internal/2d/tsdf_2d.{h,cc}: the parts the real-time matcher reads plus SetCell.
class TSDF2D : public Grid2D {
 public:
  TSDF2D(const MapLimits& limits, float truncation_distance, float max_weight,
         ValueConversionTables* conversion_tables)
      : Grid2D(limits, -truncation_distance, truncation_distance, conversion_tables),
        value_converter_(truncation_distance, max_weight, conversion_tables),
        weight_cells_(static_cast<size_t>(limits.cell_limits().num_x_cells) *
                          limits.cell_limits().num_y_cells,
                      TSDValueConverter::getUnknownWeightValue()) {}
  // What TSDF2D(const proto::Grid2D&, ...) builds (tsdf_2d.cc:35-47): bounds from the message,
  // converter from its tsdf_2d field (io/grid_proto_reader.h).
  TSDF2D(const MapLimits& limits, float min_correspondence_cost, float max_correspondence_cost,
         float truncation_distance, float max_weight, ValueConversionTables* conversion_tables)
      : Grid2D(limits, min_correspondence_cost, max_correspondence_cost, conversion_tables),
        value_converter_(truncation_distance, max_weight, conversion_tables),
        weight_cells_(static_cast<size_t>(limits.cell_limits().num_x_cells) *
                          limits.cell_limits().num_y_cells,
                      TSDValueConverter::getUnknownWeightValue()) {}
  GridType GetGridType() const override { return GridType::TSDF; }
  // tsdf_2d.cc:51-68 (the update marker is dropped at once: no FinishUpdate here)
  void SetCell(int x, int y, float tsd, float weight) {
    const size_t flat = static_cast<size_t>(limits().cell_limits().num_x_cells) * y + x;
    (*mutable_correspondence_cost_cells())[flat] = value_converter_.TSDToValue(tsd);
    weight_cells_[flat] = value_converter_.WeightToValue(weight);
  }
  const std::vector<uint16>& weight_cells() const { return weight_cells_; }
  std::vector<uint16>* mutable_weight_cells() { return &weight_cells_; }
  const TSDValueConverter& value_converter() const { return value_converter_; }

 protected:
  void Grown(const int nx, const int ny, const int x_offset, const int y_offset) override {
    const int stride = 2 * nx;
    std::vector<uint16> cells(static_cast<size_t>(4) * nx * ny, TSDValueConverter::getUnknownWeightValue());
    for (int i = 0; i < ny; ++i)
      for (int j = 0; j < nx; ++j) cells[x_offset + stride * y_offset + j + i * stride] = weight_cells_[j + i * nx];
    weight_cells_.swap(cells);
  }

 private:
  TSDValueConverter value_converter_;
  std::vector<uint16> weight_cells_;
};

namespace proto {
// mapping/proto/2d/probability_grid_range_data_inserter_options_2d.proto (fields 1-3)
class ProbabilityGridRangeDataInserterOptions2D {
 public:
  double hit_probability() const { return hit_probability_; }
  double miss_probability() const { return miss_probability_; }
  bool insert_free_space() const { return insert_free_space_; }
  void set_hit_probability(double v) { hit_probability_ = v; }
  void set_miss_probability(double v) { miss_probability_ = v; }
  void set_insert_free_space(bool v) { insert_free_space_ = v; }

 private:
  double hit_probability_ = 0., miss_probability_ = 0.;
  bool insert_free_space_ = false;
};
}  // namespace proto

// mapping/id.h:42-95
struct NodeId {
  NodeId(int trajectory_id, int node_index) : trajectory_id(trajectory_id), node_index(node_index) {}
  int trajectory_id;
  int node_index;
  bool operator==(const NodeId& o) const {
    return std::forward_as_tuple(trajectory_id, node_index) ==
           std::forward_as_tuple(o.trajectory_id, o.node_index);
  }
  bool operator<(const NodeId& o) const {
    return std::forward_as_tuple(trajectory_id, node_index) <
           std::forward_as_tuple(o.trajectory_id, o.node_index);
  }
};
struct SubmapId {
  SubmapId(int trajectory_id, int submap_index)
      : trajectory_id(trajectory_id), submap_index(submap_index) {}
  int trajectory_id;
  int submap_index;
  bool operator==(const SubmapId& o) const {
    return std::forward_as_tuple(trajectory_id, submap_index) ==
           std::forward_as_tuple(o.trajectory_id, o.submap_index);
  }
  bool operator<(const SubmapId& o) const {
    return std::forward_as_tuple(trajectory_id, submap_index) <
           std::forward_as_tuple(o.trajectory_id, o.submap_index);
  }
};

// The parts of Submap2D the constraint builder reads (mapping/2d/submap_2d.h:42-70,
// mapping/submaps.h:62-91).
class Submap2D {
 public:
  Submap2D(const transform::Rigid3d& local_pose, const Grid2D* grid)
      : local_pose_(local_pose), grid_(grid) {}
  transform::Rigid3d local_pose() const { return local_pose_; }
  const Grid2D* grid() const { return grid_; }

 private:
  transform::Rigid3d local_pose_;
  const Grid2D* grid_;
};

// mapping/trajectory_node.h:43-70 (the fields the scan-matching paths read; the
// histogram is an Eigen::VectorXf in the reference).
struct TrajectoryNode {
  struct Data {
    transform::Quaterniond gravity_alignment;
    sensor::PointCloud filtered_gravity_aligned_point_cloud;
    sensor::PointCloud high_resolution_point_cloud;
    sensor::PointCloud low_resolution_point_cloud;
    std::vector<float> rotational_scan_matcher_histogram;
  };
};

// mapping/pose_graph_interface.h:36-59
class PoseGraphInterface {
 public:
  struct Constraint {
    struct Pose {
      transform::Rigid3d zbar_ij;
      double translation_weight;
      double rotation_weight;
    };
    SubmapId submap_id;
    NodeId node_id;
    Pose pose;
    enum Tag { INTRA_SUBMAP, INTER_SUBMAP } tag;
  };
};

}  // namespace mapping
namespace metrics {
class FamilyFactory;  // cartographer/metrics/family_factory.h
}  // namespace metrics
namespace common {
namespace proto {
// common/proto/ceres_solver_options.proto (fields 1-3)
class CeresSolverOptions {
 public:
  bool use_nonmonotonic_steps() const { return use_nonmonotonic_steps_; }
  int max_num_iterations() const { return max_num_iterations_; }
  int num_threads() const { return num_threads_; }
  void set_use_nonmonotonic_steps(bool v) { use_nonmonotonic_steps_ = v; }
  void set_max_num_iterations(int v) { max_num_iterations_ = v; }
  void set_num_threads(int v) { num_threads_ = v; }

 private:
  bool use_nonmonotonic_steps_ = false;
  int max_num_iterations_ = 0;
  int num_threads_ = 1;
};
}  // namespace proto
}  // namespace common
}  // namespace cartographer

// The fields of ceres::Solver::Summary a Cartographer caller reads (ceres/solver.h); the
// solve itself runs on the device (csm_ceres_match2d_batch).
namespace ceres {
enum TerminationType { CONVERGENCE, NO_CONVERGENCE, FAILURE };
struct Solver {
  struct Summary {
    double initial_cost = 0., final_cost = 0.;
    int num_successful_steps = 0, num_unsuccessful_steps = 0;
    int num_iterations = 0;  // iterations after the initial evaluation
    TerminationType termination_type = NO_CONVERGENCE;
    std::string message;
    std::string BriefReport() const {
      return "B200 LM: iterations " + std::to_string(num_iterations) + ", initial_cost " +
             std::to_string(initial_cost) + ", final_cost " + std::to_string(final_cost) + ", " + message;
    }
    std::string FullReport() const { return BriefReport(); }
    bool IsSolutionUsable() const { return termination_type != FAILURE; }
  };
};
}  // namespace ceres

namespace cartographer {
namespace mapping {
namespace scan_matching {
namespace proto {
// mapping/proto/scan_matching/ceres_scan_matcher_options_2d.proto (fields 1, 2, 3, 9)
class CeresScanMatcherOptions2D {
 public:
  double occupied_space_weight() const { return occupied_space_weight_; }
  double translation_weight() const { return translation_weight_; }
  double rotation_weight() const { return rotation_weight_; }
  const common::proto::CeresSolverOptions& ceres_solver_options() const { return ceres_solver_options_; }
  common::proto::CeresSolverOptions* mutable_ceres_solver_options() { return &ceres_solver_options_; }
  void set_occupied_space_weight(double v) { occupied_space_weight_ = v; }
  void set_translation_weight(double v) { translation_weight_ = v; }
  void set_rotation_weight(double v) { rotation_weight_ = v; }

 private:
  double occupied_space_weight_ = 0., translation_weight_ = 0., rotation_weight_ = 0.;
  common::proto::CeresSolverOptions ceres_solver_options_;
};
// fast_correlative_scan_matcher_options_2d.proto
class FastCorrelativeScanMatcherOptions2D {
 public:
  double linear_search_window() const { return linear_search_window_; }
  double angular_search_window() const { return angular_search_window_; }
  int branch_and_bound_depth() const { return branch_and_bound_depth_; }
  void set_linear_search_window(double v) { linear_search_window_ = v; }
  void set_angular_search_window(double v) { angular_search_window_ = v; }
  void set_branch_and_bound_depth(int v) { branch_and_bound_depth_ = v; }

 private:
  double linear_search_window_ = 0.;
  double angular_search_window_ = 0.;
  int branch_and_bound_depth_ = 0;
};
// real_time_correlative_scan_matcher_options.proto
// fast_correlative_scan_matcher_options_3d.proto (fields 2, 8, 4, 9, 5, 6, 7)
class FastCorrelativeScanMatcherOptions3D {
 public:
  int branch_and_bound_depth() const { return branch_and_bound_depth_; }
  int full_resolution_depth() const { return full_resolution_depth_; }
  double min_rotational_score() const { return min_rotational_score_; }
  double min_low_resolution_score() const { return min_low_resolution_score_; }
  double linear_xy_search_window() const { return linear_xy_search_window_; }
  double linear_z_search_window() const { return linear_z_search_window_; }
  double angular_search_window() const { return angular_search_window_; }
  void set_branch_and_bound_depth(int v) { branch_and_bound_depth_ = v; }
  void set_full_resolution_depth(int v) { full_resolution_depth_ = v; }
  void set_min_rotational_score(double v) { min_rotational_score_ = v; }
  void set_min_low_resolution_score(double v) { min_low_resolution_score_ = v; }
  void set_linear_xy_search_window(double v) { linear_xy_search_window_ = v; }
  void set_linear_z_search_window(double v) { linear_z_search_window_ = v; }
  void set_angular_search_window(double v) { angular_search_window_ = v; }

 private:
  int branch_and_bound_depth_ = 0, full_resolution_depth_ = 0;
  double min_rotational_score_ = 0., min_low_resolution_score_ = 0.;
  double linear_xy_search_window_ = 0., linear_z_search_window_ = 0., angular_search_window_ = 0.;
};
class RealTimeCorrelativeScanMatcherOptions {
 public:
  double linear_search_window() const { return linear_search_window_; }
  double angular_search_window() const { return angular_search_window_; }
  double translation_delta_cost_weight() const { return translation_delta_cost_weight_; }
  double rotation_delta_cost_weight() const { return rotation_delta_cost_weight_; }
  void set_linear_search_window(double v) { linear_search_window_ = v; }
  void set_angular_search_window(double v) { angular_search_window_ = v; }
  void set_translation_delta_cost_weight(double v) { translation_delta_cost_weight_ = v; }
  void set_rotation_delta_cost_weight(double v) { rotation_delta_cost_weight_ = v; }

 private:
  double linear_search_window_ = 0., angular_search_window_ = 0.;
  double translation_delta_cost_weight_ = 0., rotation_delta_cost_weight_ = 0.;
};
}  // namespace proto
}  // namespace scan_matching
namespace constraints {
namespace proto {
// mapping/proto/pose_graph/constraint_builder_options.proto
class ConstraintBuilderOptions {
 public:
  double sampling_ratio() const { return sampling_ratio_; }
  double max_constraint_distance() const { return max_constraint_distance_; }
  double min_score() const { return min_score_; }
  double global_localization_min_score() const { return global_localization_min_score_; }
  double loop_closure_translation_weight() const { return loop_closure_translation_weight_; }
  double loop_closure_rotation_weight() const { return loop_closure_rotation_weight_; }
  bool log_matches() const { return log_matches_; }
  const scan_matching::proto::FastCorrelativeScanMatcherOptions2D&
  fast_correlative_scan_matcher_options() const {
    return fast_correlative_scan_matcher_options_;
  }
  scan_matching::proto::FastCorrelativeScanMatcherOptions2D*
  mutable_fast_correlative_scan_matcher_options() {
    return &fast_correlative_scan_matcher_options_;
  }
  const scan_matching::proto::FastCorrelativeScanMatcherOptions3D&
  fast_correlative_scan_matcher_options_3d() const {
    return fast_correlative_scan_matcher_options_3d_;
  }
  scan_matching::proto::FastCorrelativeScanMatcherOptions3D*
  mutable_fast_correlative_scan_matcher_options_3d() {
    return &fast_correlative_scan_matcher_options_3d_;
  }
  // field 11; an occupied_space_weight of 0 (message not filled in) disables the refinement
  const scan_matching::proto::CeresScanMatcherOptions2D& ceres_scan_matcher_options() const {
    return ceres_scan_matcher_options_;
  }
  scan_matching::proto::CeresScanMatcherOptions2D* mutable_ceres_scan_matcher_options() {
    return &ceres_scan_matcher_options_;
  }
  void set_sampling_ratio(double v) { sampling_ratio_ = v; }
  void set_max_constraint_distance(double v) { max_constraint_distance_ = v; }
  void set_min_score(double v) { min_score_ = v; }
  void set_global_localization_min_score(double v) { global_localization_min_score_ = v; }
  void set_loop_closure_translation_weight(double v) { loop_closure_translation_weight_ = v; }
  void set_loop_closure_rotation_weight(double v) { loop_closure_rotation_weight_ = v; }
  void set_log_matches(bool v) { log_matches_ = v; }

 private:
  double sampling_ratio_ = 0., max_constraint_distance_ = 0., min_score_ = 0.;
  double global_localization_min_score_ = 0.;
  double loop_closure_translation_weight_ = 0., loop_closure_rotation_weight_ = 0.;
  bool log_matches_ = false;
  scan_matching::proto::FastCorrelativeScanMatcherOptions2D fast_correlative_scan_matcher_options_;
  scan_matching::proto::FastCorrelativeScanMatcherOptions3D fast_correlative_scan_matcher_options_3d_;
  scan_matching::proto::CeresScanMatcherOptions2D ceres_scan_matcher_options_;
};
}  // namespace proto
}  // namespace constraints
}  // namespace mapping
}  // namespace cartographer

#endif  // CSM_HOST_CARTOGRAPHER_COMPAT_H_
