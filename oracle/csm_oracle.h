// TEST INFRASTRUCTURE ONLY -- CPU oracle for the correlative-scan-matching hot path.
//
// This header restates, on the CPU and without any third-party dependency, the
// algorithm of juwangvsu/cartographer-1's 2D correlative scan matchers so that the
// CUDA path of this repository can be checked bit-for-bit against it.  Nothing in
// the product path (cartographer-1_b200/) may include, link or execute this code;
// only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference legs do.
//
// Parity status: the reference's hot-path translation units need Eigen, glog,
// abseil, protobuf, Lua and Ceres, none of which exist in this image, so the
// reference itself could not be compiled here ("unbuildable").  The restatement is
// pinned against every known-answer test the reference holds for this path
// (oracle/oracle_tests.cc ports them; see each TEST for the reference file:line).
// Those reference tests are tolerance tests; nothing in the reference pins the
// bit-level winner (scan_index, x_offset, y_offset) or tie order, so bit-level
// parity is "unpinned" beyond this restatement.
//
// Third-party arithmetic restated here (absent from /root/reference):
//   * Eigen 3.3 (bazel pin f3a22f35b044): Quaternionf(AngleAxisf), quaternion *
//     vector (QuaternionBase::_transformVector), Vector2f::norm, Affine2f * v,
//     Rotation2D composition.  All fp32/fp64, no FMA contraction (the reference is
//     built -O3 for baseline x86-64; this file must be compiled -ffp-contract=off).
//   * glibc libm cosf/sinf/acos/ceil/lround/hypot/exp and libstdc++ std::sort
//     (introsort; its tie permutation is part of the reference's observable result).
//
// Citations are relative to /root/reference/cartographer/.
#ifndef ORACLE_CSM_ORACLE_H_
#define ORACLE_CSM_ORACLE_H_

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <deque>
#include <functional>
#include <limits>
#include <memory>
#include <vector>

namespace oracle {

#define ORACLE_CHECK(cond)                                                      \
  do {                                                                          \
    if (!(cond)) {                                                              \
      std::fprintf(stderr, "ORACLE_CHECK failed: %s (%s:%d)\n", #cond, __FILE__, \
                   __LINE__);                                                   \
      std::abort();                                                             \
    }                                                                           \
  } while (0)

// common/port.h:40-42
inline int RoundToInt(const float x) { return static_cast<int>(std::lround(x)); }
inline int RoundToInt(const double x) { return static_cast<int>(std::lround(x)); }

// common/math.h:31-40
template <typename T>
inline T Clamp(const T v, const T lo, const T hi) {
  if (v > hi) return hi;
  if (v < lo) return lo;
  return v;
}

struct Array2i {
  int x = 0, y = 0;
};
struct Vec3f {
  float x = 0.f, y = 0.f, z = 0.f;
};
typedef std::vector<Vec3f> PointCloud;  // positions only; sensor/point_cloud.h:33-92

// transform/rigid_transform.h:34-103 (Rigid2d); the rotation is kept as its angle
// exactly like Eigen::Rotation2Dd does.
struct Rigid2d {
  double tx = 0., ty = 0., angle = 0.;
};

// ---------------------------------------------------------------------------
// Eigen restatements.
// ---------------------------------------------------------------------------
// Quaternionf(AngleAxisf(angle, UnitZ)): w = cos(angle/2), vec = sin(angle/2)*axis.
struct Quatf {
  float w, x, y, z;
};
inline Quatf QuatFromYaw(const float angle) {
  const float half = 0.5f * angle;
  const float s = std::sin(half);  // sinf
  Quatf q;
  q.w = std::cos(half);  // cosf
  q.x = s * 0.f;
  q.y = s * 0.f;
  q.z = s * 1.f;
  return q;
}
// Eigen 3.3 QuaternionBase::_transformVector: uv = vec x v; uv += uv;
// result = v + w*uv + vec x uv   (evaluated as (v + w*uv) + cross).
inline Vec3f QuatRotate(const Quatf& q, const Vec3f& v) {
  Vec3f uv;
  uv.x = q.y * v.z - q.z * v.y;
  uv.y = q.z * v.x - q.x * v.z;
  uv.z = q.x * v.y - q.y * v.x;
  uv.x += uv.x;
  uv.y += uv.y;
  uv.z += uv.z;
  Vec3f c;
  c.x = q.y * uv.z - q.z * uv.y;
  c.y = q.z * uv.x - q.x * uv.z;
  c.z = q.x * uv.y - q.y * uv.x;
  Vec3f r;
  r.x = (v.x + q.w * uv.x) + c.x;
  r.y = (v.y + q.w * uv.y) + c.y;
  r.z = (v.z + q.w * uv.z) + c.z;
  return r;
}
// sensor::TransformPointCloud with Rigid3f::Rotation(AngleAxisf(angle, UnitZ))
// (sensor/point_cloud.cc:56-64, transform/rigid_transform.h:129-131,192-196):
// rotation * point + Vector3f::Zero().
inline PointCloud RotatePointCloudYaw(const PointCloud& cloud, const float angle) {
  const Quatf q = QuatFromYaw(angle);
  PointCloud out;
  out.reserve(cloud.size());
  for (const Vec3f& p : cloud) {
    Vec3f r = QuatRotate(q, p);
    r.x = r.x + 0.f;
    r.y = r.y + 0.f;
    r.z = r.z + 0.f;
    out.push_back(r);
  }
  return out;
}

// ---------------------------------------------------------------------------
// Value tables.  mapping/value_conversion_tables.cc:29-67 and
// mapping/probability_values.{h,cc}.
// ---------------------------------------------------------------------------
constexpr float kMinProbability = 0.1f;                         // probability_values.h:56
constexpr float kMaxProbability = 1.f - kMinProbability;        // :57
constexpr float kMinCorrespondenceCost = 1.f - kMaxProbability; // :58
constexpr float kMaxCorrespondenceCost = 1.f - kMinProbability; // :59
constexpr uint16_t kUnknownValue = 0;
constexpr uint16_t kUpdateMarker = 1u << 15;

inline float SlowValueToBoundedFloat(const uint16_t value, const uint16_t unknown_value,
                                     const float unknown_result, const float lower,
                                     const float upper) {
  if (value == unknown_value) return unknown_result;
  const float scale = (upper - lower) / 32766.f;
  return value * scale + (lower - scale);
}
// 65536 entries; the update-marker bit is masked off (value_conversion_tables.cc:41-52).
inline std::vector<float> MakeConversionTable(const float unknown_result, const float lower,
                                              const float upper) {
  std::vector<float> t;
  t.reserve(65536);
  for (uint32_t v = 0; v != 65536; ++v) {
    t.push_back(SlowValueToBoundedFloat(static_cast<uint16_t>(v) & ~kUpdateMarker, 0,
                                        unknown_result, lower, upper));
  }
  return t;
}
inline const std::vector<float>& ValueToCorrespondenceCostTable() {
  static const std::vector<float> t =
      MakeConversionTable(kMaxCorrespondenceCost, kMinCorrespondenceCost, kMaxCorrespondenceCost);
  return t;
}
inline const std::vector<float>& ValueToProbabilityTable() {
  static const std::vector<float> t =
      MakeConversionTable(kMinProbability, kMinProbability, kMaxProbability);
  return t;
}
// probability_values.h:32-46
inline uint16_t BoundedFloatToValue(const float v, const float lower, const float upper) {
  const int value =
      RoundToInt((Clamp(v, lower, upper) - lower) * (32766.f / (upper - lower))) + 1;
  return static_cast<uint16_t>(value);
}
inline uint16_t CorrespondenceCostToValue(const float c) {
  return BoundedFloatToValue(c, kMinCorrespondenceCost, kMaxCorrespondenceCost);
}
inline uint16_t ProbabilityToValue(const float p) {
  return BoundedFloatToValue(p, kMinProbability, kMaxProbability);
}
inline float Odds(const float p) { return p / (1.f - p); }
inline float ProbabilityFromOdds(const float o) { return o / (o + 1.f); }
// probability_values.cc:89-105
inline std::vector<uint16_t> ComputeLookupTableToApplyCorrespondenceCostOdds(const float odds) {
  std::vector<uint16_t> result;
  result.reserve(32768);
  result.push_back(CorrespondenceCostToValue(1.f - ProbabilityFromOdds(odds)) + kUpdateMarker);
  const std::vector<float>& cc = ValueToCorrespondenceCostTable();
  for (int cell = 1; cell != 32768; ++cell) {
    result.push_back(
        CorrespondenceCostToValue(1.f - ProbabilityFromOdds(odds * Odds(1.f - cc[cell]))) +
        kUpdateMarker);
  }
  return result;
}

// ---------------------------------------------------------------------------
// MapLimits (mapping/2d/map_limits.h:39-98) and the probability grid
// (mapping/2d/grid_2d.{h,cc}, mapping/2d/probability_grid.cc).
// ---------------------------------------------------------------------------
struct CellLimits {
  int num_x_cells = 0, num_y_cells = 0;
};
struct MapLimits {
  double resolution = 0.;
  double max_x = 0., max_y = 0.;
  CellLimits cell_limits;

  // map_limits.h:69-76.  The cell .x comes from the world y coordinate.
  Array2i GetCellIndex(const float px, const float py) const {
    Array2i c;
    c.x = RoundToInt((max_y - py) / resolution - 0.5);
    c.y = RoundToInt((max_x - px) / resolution - 0.5);
    return c;
  }
  bool Contains(const Array2i& c) const {
    return c.x >= 0 && c.y >= 0 && c.x < cell_limits.num_x_cells && c.y < cell_limits.num_y_cells;
  }
};

class ProbabilityGrid {
 public:
  ProbabilityGrid(const MapLimits& limits)
      : limits_(limits),
        cells_(static_cast<size_t>(limits.cell_limits.num_x_cells) * limits.cell_limits.num_y_cells,
               kUnknownValue),
        table_(&ValueToCorrespondenceCostTable()) {}
  // Adopts an existing cell array (flat = num_x_cells * y + x, grid_2d.h:113-116).
  ProbabilityGrid(const MapLimits& limits, const uint16_t* cells)
      : limits_(limits),
        cells_(cells, cells + static_cast<size_t>(limits.cell_limits.num_x_cells) *
                                  limits.cell_limits.num_y_cells),
        table_(&ValueToCorrespondenceCostTable()) {}

  const MapLimits& limits() const { return limits_; }
  const std::vector<uint16_t>& cells() const { return cells_; }
  float GetMinCorrespondenceCost() const { return kMinCorrespondenceCost; }
  float GetMaxCorrespondenceCost() const { return kMaxCorrespondenceCost; }

  int ToFlatIndex(const Array2i& c) const {
    ORACLE_CHECK(limits_.Contains(c));
    return limits_.cell_limits.num_x_cells * c.y + c.x;
  }
  // grid_2d.h:53-57
  float GetCorrespondenceCost(const Array2i& c) const {
    if (!limits_.Contains(c)) return kMaxCorrespondenceCost;
    return (*table_)[cells_[ToFlatIndex(c)]];
  }
  // probability_grid.cc:78-82
  float GetProbability(const Array2i& c) const {
    if (!limits_.Contains(c)) return kMinProbability;
    return 1.f - ValueToCorrespondenceCostTable()[cells_[ToFlatIndex(c)]];
  }
  // probability_grid.cc:40-48
  void SetProbability(const Array2i& c, const float probability) {
    uint16_t& cell = cells_[ToFlatIndex(c)];
    ORACLE_CHECK(cell == kUnknownValue);
    cell = CorrespondenceCostToValue(1.f - probability);
  }
  // probability_grid.cc:57-71
  bool ApplyLookupTable(const Array2i& c, const std::vector<uint16_t>& table) {
    const int flat = ToFlatIndex(c);
    uint16_t* cell = &cells_[flat];
    if (*cell >= kUpdateMarker) return false;
    update_indices_.push_back(flat);
    *cell = table[*cell];
    return true;
  }
  // grid_2d.cc:98-106
  void FinishUpdate() {
    while (!update_indices_.empty()) {
      cells_[update_indices_.back()] -= kUpdateMarker;
      update_indices_.pop_back();
    }
  }
  // grid_2d.cc:131-166
  void GrowLimits(const float px, const float py) {
    ORACLE_CHECK(update_indices_.empty());
    while (!limits_.Contains(limits_.GetCellIndex(px, py))) {
      const int nx = limits_.cell_limits.num_x_cells, ny = limits_.cell_limits.num_y_cells;
      const int x_offset = nx / 2, y_offset = ny / 2;
      MapLimits grown;
      grown.resolution = limits_.resolution;
      grown.max_x = limits_.max_x + limits_.resolution * y_offset;
      grown.max_y = limits_.max_y + limits_.resolution * x_offset;
      grown.cell_limits.num_x_cells = 2 * nx;
      grown.cell_limits.num_y_cells = 2 * ny;
      const int stride = grown.cell_limits.num_x_cells;
      std::vector<uint16_t> cells(static_cast<size_t>(4) * nx * ny, kUnknownValue);
      for (int i = 0; i < ny; ++i)
        for (int j = 0; j < nx; ++j)
          cells[x_offset + stride * y_offset + j + i * stride] = cells_[j + i * nx];
      cells_.swap(cells);
      limits_ = grown;
    }
  }

 private:
  MapLimits limits_;
  std::vector<uint16_t> cells_;
  std::vector<int> update_indices_;
  const std::vector<float>* table_;
};

// ---------------------------------------------------------------------------
// TSDF2D: the second Grid2D flavour the real-time matcher scores against
// (mapping/internal/2d/tsdf_2d.{h,cc}, tsd_value_converter.{h,cc}).  Two uint16 arrays
// (truncated signed distance, weight), each with its own value table.
// ---------------------------------------------------------------------------
class TSDF2D {
 public:
  // tsdf_2d.cc:24-33, tsd_value_converter.cc:22-32
  TSDF2D(const MapLimits& limits, const float truncation_distance, const float max_weight)
      : limits_(limits),
        max_tsd_(truncation_distance),
        min_tsd_(-truncation_distance),
        max_weight_(max_weight),
        tsd_resolution_(32766.f / (max_tsd_ - min_tsd_)),
        weight_resolution_(32766.f / (max_weight_ - 0.f)),
        tsd_cells_(static_cast<size_t>(limits.cell_limits.num_x_cells) * limits.cell_limits.num_y_cells,
                   kUnknownValue),
        weight_cells_(tsd_cells_.size(), kUnknownValue),
        value_to_tsd_(MakeConversionTable(min_tsd_, min_tsd_, max_tsd_)),
        value_to_weight_(MakeConversionTable(0.f, 0.f, max_weight_)),
        value_to_cost_(MakeConversionTable(max_tsd_, min_tsd_, max_tsd_)) {}
  // Adopts existing cell arrays (flat = num_x_cells * y + x).
  TSDF2D(const MapLimits& limits, const float truncation_distance, const float max_weight,
         const uint16_t* tsd_cells, const uint16_t* weight_cells)
      : TSDF2D(limits, truncation_distance, max_weight) {
    tsd_cells_.assign(tsd_cells, tsd_cells + tsd_cells_.size());
    weight_cells_.assign(weight_cells, weight_cells + weight_cells_.size());
  }

  const MapLimits& limits() const { return limits_; }
  const std::vector<uint16_t>& tsd_cells() const { return tsd_cells_; }
  const std::vector<uint16_t>& weight_cells() const { return weight_cells_; }
  const std::vector<float>& value_to_tsd() const { return value_to_tsd_; }
  const std::vector<float>& value_to_weight() const { return value_to_weight_; }
  // Grid2D's max_correspondence_cost_ of a TSDF2D is the truncation distance (tsdf_2d.cc:26).
  float GetMaxCorrespondenceCost() const { return max_tsd_; }
  // Grid2D(limits, -truncation_distance, truncation_distance, tables) (tsdf_2d.cc:23-27).
  float GetMinCorrespondenceCost() const { return min_tsd_; }
  float GetMinTSD() const { return min_tsd_; }
  // Grid2D::GetCorrespondenceCost (grid_2d.h:53-57) of a TSDF2D: the base class's own table is
  // GetConversionTable(max_cost, min_cost, max_cost) (grid_2d.cc:66-68), i.e. unknown reads
  // +truncation_distance -- not TSDValueConverter's table, where unknown reads min_tsd.
  float GetCorrespondenceCost(const Array2i& c) const {
    if (!limits_.Contains(c)) return max_tsd_;
    return value_to_cost_[tsd_cells_[limits_.cell_limits.num_x_cells * c.y + c.x]];
  }

  // tsd_value_converter.h:38-56
  uint16_t TSDToValue(const float tsd) const {
    return static_cast<uint16_t>(RoundToInt((Clamp(tsd, min_tsd_, max_tsd_) - min_tsd_) * tsd_resolution_) + 1);
  }
  uint16_t WeightToValue(const float weight) const {
    return static_cast<uint16_t>(RoundToInt((Clamp(weight, 0.f, max_weight_) - 0.f) * weight_resolution_) + 1);
  }
  // tsdf_2d.cc:51-68 without the update marker (cells are painted once, then "finished").
  void SetCell(const Array2i& c, const float tsd, const float weight) {
    ORACLE_CHECK(limits_.Contains(c));
    const int flat = limits_.cell_limits.num_x_cells * c.y + c.x;
    tsd_cells_[flat] = TSDToValue(tsd);
    weight_cells_[flat] = WeightToValue(weight);
  }
  // tsdf_2d.cc:44-48, 51-68 with the update marker, and Grid2D::FinishUpdate (grid_2d.cc:98-106):
  // what TSDFRangeDataInserter2D uses (csm_tsdf_inserter.h).
  bool CellIsUpdated(const Array2i& c) const {
    ORACLE_CHECK(limits_.Contains(c));
    return tsd_cells_[limits_.cell_limits.num_x_cells * c.y + c.x] >= kUpdateMarker;
  }
  void SetCellMarked(const Array2i& c, const float tsd, const float weight) {
    ORACLE_CHECK(limits_.Contains(c));
    const int flat = limits_.cell_limits.num_x_cells * c.y + c.x;
    if (tsd_cells_[flat] >= kUpdateMarker) return;
    update_indices_.push_back(flat);
    tsd_cells_[flat] = static_cast<uint16_t>(TSDToValue(tsd) + kUpdateMarker);
    weight_cells_[flat] = WeightToValue(weight);
  }
  void FinishUpdate() {
    while (!update_indices_.empty()) {
      tsd_cells_[update_indices_.back()] -= kUpdateMarker;
      update_indices_.pop_back();
    }
  }
  // tsdf_2d.cc:100-105 on grid_2d.cc:131-166: both cell arrays, unknown value 0 for each.
  void GrowLimits(const float px, const float py) {
    ORACLE_CHECK(update_indices_.empty());
    while (!limits_.Contains(limits_.GetCellIndex(px, py))) {
      const int nx = limits_.cell_limits.num_x_cells, ny = limits_.cell_limits.num_y_cells;
      const int x_offset = nx / 2, y_offset = ny / 2;
      MapLimits grown;
      grown.resolution = limits_.resolution;
      grown.max_x = limits_.max_x + limits_.resolution * y_offset;
      grown.max_y = limits_.max_y + limits_.resolution * x_offset;
      grown.cell_limits.num_x_cells = 2 * nx;
      grown.cell_limits.num_y_cells = 2 * ny;
      const int stride = grown.cell_limits.num_x_cells;
      for (std::vector<uint16_t>* grid : {&tsd_cells_, &weight_cells_}) {
        std::vector<uint16_t> cells(static_cast<size_t>(4) * nx * ny, kUnknownValue);
        for (int i = 0; i < ny; ++i)
          for (int j = 0; j < nx; ++j) cells[x_offset + stride * y_offset + j + i * stride] = (*grid)[j + i * nx];
        grid->swap(cells);
      }
      limits_ = grown;
    }
  }
  float max_weight() const { return max_weight_; }
  float truncation_distance() const { return max_tsd_; }

  // tsdf_2d.cc:88-98
  std::pair<float, float> GetTSDAndWeight(const Array2i& c) const {
    if (limits_.Contains(c)) {
      const int flat = limits_.cell_limits.num_x_cells * c.y + c.x;
      return std::make_pair(value_to_tsd_[tsd_cells_[flat]], value_to_weight_[weight_cells_[flat]]);
    }
    return std::make_pair(min_tsd_, 0.f);
  }

 private:
  MapLimits limits_;
  float max_tsd_, min_tsd_, max_weight_, tsd_resolution_, weight_resolution_;
  std::vector<uint16_t> tsd_cells_, weight_cells_;
  std::vector<int> update_indices_;
  std::vector<float> value_to_tsd_, value_to_weight_, value_to_cost_;
};

// ---------------------------------------------------------------------------
// Range-data insertion: only needed to build the fixtures of the reference's
// FastCorrelativeScanMatcherTest (mapping/2d/probability_grid_range_data_inserter_2d.cc:30-143,
// mapping/internal/2d/ray_to_pixel_mask.cc:25-170).
// ---------------------------------------------------------------------------
inline bool SameCell(const Array2i& a, const Array2i& b) { return a.x == b.x && a.y == b.y; }

inline std::vector<Array2i> RayToPixelMask(const Array2i& b, const Array2i& e, const int scale) {
  if (b.x > e.x) return RayToPixelMask(e, b, scale);
  ORACLE_CHECK(b.x >= 0 && b.y >= 0 && e.y >= 0);
  std::vector<Array2i> mask;
  auto push_if_new = [&mask](const Array2i& c) {
    if (!SameCell(mask.back(), c)) mask.push_back(c);
  };
  if (b.x / scale == e.x / scale) {  // vertical in full pixels
    Array2i cur;
    cur.x = b.x / scale;
    cur.y = std::min(b.y, e.y) / scale;
    mask.push_back(cur);
    const int end_y = std::max(b.y, e.y) / scale;
    for (; cur.y <= end_y; ++cur.y) push_if_new(cur);
    return mask;
  }
  const int64_t dx = e.x - b.x;
  const int64_t dy = e.y - b.y;
  const int64_t denom = 2 * static_cast<int64_t>(scale) * dx;
  Array2i cur;
  cur.x = b.x / scale;
  cur.y = b.y / scale;
  mask.push_back(cur);
  int64_t sub_y = (2 * (b.y % scale) + 1) * dx;
  const int first_pixel = 2 * scale - 2 * (b.x % scale) - 1;
  const int last_pixel = 2 * (e.x % scale) + 1;
  const int end_x = std::max(b.x, e.x) / scale;
  sub_y += dy * first_pixel;
  const int dir = dy > 0 ? 1 : -1;
  // Ascending rays step while sub_y exceeds the denominator, the others while it
  // is negative; both variants of the reference loop are folded into one here.
  auto over = [&](int64_t s) { return dir > 0 ? s > denom : s < 0; };
  auto on_edge = [&](int64_t s) { return dir > 0 ? s == denom : s == 0; };
  while (true) {
    push_if_new(cur);
    while (over(sub_y)) {
      sub_y -= dir * denom;
      cur.y += dir;
      push_if_new(cur);
    }
    ++cur.x;
    if (on_edge(sub_y)) {
      sub_y -= dir * denom;
      cur.y += dir;
    }
    if (cur.x == end_x) break;
    sub_y += dy * 2 * scale;
  }
  sub_y += dy * last_pixel;
  push_if_new(cur);
  while (over(sub_y)) {
    sub_y -= dir * denom;
    cur.y += dir;
    push_if_new(cur);
  }
  ORACLE_CHECK(!on_edge(sub_y));
  ORACLE_CHECK(cur.y == e.y / scale);
  return mask;
}

struct RangeData {
  Vec3f origin;
  PointCloud returns;
  PointCloud misses;
};

class ProbabilityGridRangeDataInserter2D {
 public:
  ProbabilityGridRangeDataInserter2D(const double hit_probability, const double miss_probability,
                                     const bool insert_free_space)
      : insert_free_space_(insert_free_space),
        hit_table_(ComputeLookupTableToApplyCorrespondenceCostOdds(
            Odds(static_cast<float>(hit_probability)))),
        miss_table_(ComputeLookupTableToApplyCorrespondenceCostOdds(
            Odds(static_cast<float>(miss_probability)))) {}

  void Insert(const RangeData& rd, ProbabilityGrid* grid) const {
    constexpr int kSubpixelScale = 1000;
    // GrowAsNeeded: bounding box of origin + returns + misses, padded by 1e-6.
    float min_x = rd.origin.x, max_x = rd.origin.x, min_y = rd.origin.y, max_y = rd.origin.y;
    auto extend = [&](const Vec3f& p) {
      min_x = std::min(min_x, p.x);
      max_x = std::max(max_x, p.x);
      min_y = std::min(min_y, p.y);
      max_y = std::max(max_y, p.y);
    };
    for (const Vec3f& p : rd.returns) extend(p);
    for (const Vec3f& p : rd.misses) extend(p);
    constexpr float kPadding = 1e-6f;
    grid->GrowLimits(min_x - kPadding * 1.f, min_y - kPadding * 1.f);
    grid->GrowLimits(max_x + kPadding * 1.f, max_y + kPadding * 1.f);

    const MapLimits& limits = grid->limits();
    MapLimits super;
    super.resolution = limits.resolution / kSubpixelScale;
    super.max_x = limits.max_x;
    super.max_y = limits.max_y;
    super.cell_limits.num_x_cells = limits.cell_limits.num_x_cells * kSubpixelScale;
    super.cell_limits.num_y_cells = limits.cell_limits.num_y_cells * kSubpixelScale;
    const Array2i begin = super.GetCellIndex(rd.origin.x, rd.origin.y);
    std::vector<Array2i> ends;
    ends.reserve(rd.returns.size());
    for (const Vec3f& hit : rd.returns) {
      ends.push_back(super.GetCellIndex(hit.x, hit.y));
      Array2i c;
      c.x = ends.back().x / kSubpixelScale;
      c.y = ends.back().y / kSubpixelScale;
      grid->ApplyLookupTable(c, hit_table_);
    }
    if (insert_free_space_) {
      for (const Array2i& end : ends)
        for (const Array2i& c : RayToPixelMask(begin, end, kSubpixelScale))
          grid->ApplyLookupTable(c, miss_table_);
      for (const Vec3f& miss : rd.misses)
        for (const Array2i& c :
             RayToPixelMask(begin, super.GetCellIndex(miss.x, miss.y), kSubpixelScale))
          grid->ApplyLookupTable(c, miss_table_);
    }
    grid->FinishUpdate();
  }

 private:
  bool insert_free_space_;
  std::vector<uint16_t> hit_table_, miss_table_;
};

// ---------------------------------------------------------------------------
// Search parameters, rotated scans, discretisation, candidates.
// mapping/internal/2d/scan_matching/correlative_scan_matcher_2d.{h,cc}
// ---------------------------------------------------------------------------
typedef std::vector<Array2i> DiscreteScan2D;

struct LinearBounds {
  int min_x, max_x, min_y, max_y;
};

struct SearchParameters {
  // correlative_scan_matcher_2d.cc:27-55
  SearchParameters(const double linear_search_window, const double angular_search_window,
                   const PointCloud& cloud, const double res)
      : resolution(res) {
    float max_scan_range = 3.f * res;
    for (const Vec3f& p : cloud) {
      const float range = std::sqrt(p.x * p.x + p.y * p.y);  // head<2>().norm()
      max_scan_range = std::max(range, max_scan_range);
    }
    const double kSafetyMargin = 1. - 1e-3;
    // common::Pow2<float>(max_scan_range) is an fp32 product (common/math.h:43-52).
    const float r2 = max_scan_range * (max_scan_range * 1.f);
    angular_perturbation_step_size =
        kSafetyMargin * std::acos(1. - (res * (res * 1.)) / (2. * r2));
    num_angular_perturbations =
        static_cast<int>(std::ceil(angular_search_window / angular_perturbation_step_size));
    num_scans = 2 * num_angular_perturbations + 1;
    const int n = static_cast<int>(std::ceil(linear_search_window / res));
    linear_bounds.assign(num_scans, LinearBounds{-n, n, -n, n});
  }
  // correlative_scan_matcher_2d.cc:57-71 ("for testing")
  SearchParameters(const int num_linear, const int num_angular, const double step, const double res)
      : num_angular_perturbations(num_angular),
        angular_perturbation_step_size(step),
        resolution(res),
        num_scans(2 * num_angular + 1) {
    linear_bounds.assign(num_scans, LinearBounds{-num_linear, num_linear, -num_linear, num_linear});
  }
  // correlative_scan_matcher_2d.cc:73-91
  void ShrinkToFit(const std::vector<DiscreteScan2D>& scans, const CellLimits& cl) {
    ORACLE_CHECK(static_cast<int>(scans.size()) == num_scans);
    for (int i = 0; i != num_scans; ++i) {
      int min_bx = 0, min_by = 0, max_bx = 0, max_by = 0;
      for (const Array2i& c : scans[i]) {
        min_bx = std::min(min_bx, -c.x);
        min_by = std::min(min_by, -c.y);
        max_bx = std::max(max_bx, cl.num_x_cells - 1 - c.x);
        max_by = std::max(max_by, cl.num_y_cells - 1 - c.y);
      }
      LinearBounds& lb = linear_bounds[i];
      lb.min_x = std::max(lb.min_x, min_bx);
      lb.max_x = std::min(lb.max_x, max_bx);
      lb.min_y = std::max(lb.min_y, min_by);
      lb.max_y = std::min(lb.max_y, max_by);
    }
  }

  int num_angular_perturbations;
  double angular_perturbation_step_size;
  double resolution;
  int num_scans;
  std::vector<LinearBounds> linear_bounds;
};

// correlative_scan_matcher_2d.cc:93-109.  delta_theta is a double accumulated by
// repeated addition and narrowed to float for AngleAxisf.
inline std::vector<PointCloud> GenerateRotatedScans(const PointCloud& cloud,
                                                    const SearchParameters& sp) {
  std::vector<PointCloud> scans;
  scans.reserve(sp.num_scans);
  double delta_theta = -sp.num_angular_perturbations * sp.angular_perturbation_step_size;
  for (int s = 0; s < sp.num_scans; ++s, delta_theta += sp.angular_perturbation_step_size) {
    scans.push_back(RotatePointCloudYaw(cloud, static_cast<float>(delta_theta)));
  }
  return scans;
}

// correlative_scan_matcher_2d.cc:111-127.  Affine2f(Translation2f) * p.head<2>()
// is t + (1*x + 0*y) per component in fp32.
inline std::vector<DiscreteScan2D> DiscretizeScans(const MapLimits& limits,
                                                   const std::vector<PointCloud>& scans,
                                                   const float tx, const float ty) {
  std::vector<DiscreteScan2D> out;
  out.reserve(scans.size());
  for (const PointCloud& scan : scans) {
    out.emplace_back();
    out.back().reserve(scan.size());
    for (const Vec3f& p : scan) {
      const float wx = tx + (1.f * p.x + 0.f * p.y);
      const float wy = ty + (0.f * p.x + 1.f * p.y);
      out.back().push_back(limits.GetCellIndex(wx, wy));
    }
  }
  return out;
}

// correlative_scan_matcher_2d.h:74-103
struct Candidate2D {
  Candidate2D(const int scan, const int xo, const int yo, const SearchParameters& sp)
      : scan_index(scan),
        x_index_offset(xo),
        y_index_offset(yo),
        x(-y_index_offset * sp.resolution),
        y(-x_index_offset * sp.resolution),
        orientation((scan_index - sp.num_angular_perturbations) *
                    sp.angular_perturbation_step_size) {}
  int scan_index = 0;
  int x_index_offset = 0;
  int y_index_offset = 0;
  double x = 0., y = 0., orientation = 0.;
  float score = 0.f;
  bool operator<(const Candidate2D& o) const { return score < o.score; }
  bool operator>(const Candidate2D& o) const { return score > o.score; }
};

// ---------------------------------------------------------------------------
// Precomputation grids.  fast_correlative_scan_matcher_2d.{h,cc}
// ---------------------------------------------------------------------------
// fast_correlative_scan_matcher_2d.cc:41-74: monotone deque of candidates for the
// window maximum.
class SlidingWindowMaximum {
 public:
  void AddValue(const float v) {
    while (!q_.empty() && v > q_.back()) q_.pop_back();
    q_.push_back(v);
  }
  void RemoveValue(const float v) {
    if (v == q_.front()) q_.pop_front();
  }
  float GetMaximum() const { return q_.front(); }
  bool empty() const { return q_.empty(); }

 private:
  std::deque<float> q_;
};

class PrecomputationGrid2D {
 public:
  // fast_correlative_scan_matcher_2d.cc:91-161
  // `Grid` is any Grid2D flavour (ProbabilityGrid or TSDF2D): the reference takes the base class
  // and only calls GetCorrespondenceCost / GetMin- / GetMaxCorrespondenceCost on it.
  template <typename Grid>
  PrecomputationGrid2D(const Grid& grid, const CellLimits& limits, const int width,
                       std::vector<float>* scratch)
      : offset_x_(-width + 1),
        offset_y_(-width + 1),
        wide_x_(limits.num_x_cells + width - 1),
        wide_y_(limits.num_y_cells + width - 1),
        min_score_(1.f - grid.GetMaxCorrespondenceCost()),
        max_score_(1.f - grid.GetMinCorrespondenceCost()),
        cells_(static_cast<size_t>(wide_x_) * wide_y_) {
    ORACLE_CHECK(width >= 1 && limits.num_x_cells >= 1 && limits.num_y_cells >= 1);
    const int stride = wide_x_;
    const int nx = limits.num_x_cells, ny = limits.num_y_cells;
    auto value_at = [&grid](int x, int y) {
      Array2i c;
      c.x = x;
      c.y = y;
      return 1.f - std::abs(grid.GetCorrespondenceCost(c));
    };
    std::vector<float>& inter = *scratch;
    inter.resize(static_cast<size_t>(wide_x_) * ny);
    // Row pass: inter[x0 + width - 1][y] = max over x0 <= x < x0 + width.
    for (int y = 0; y != ny; ++y) {
      SlidingWindowMaximum win;
      win.AddValue(value_at(0, y));
      for (int x = -width + 1; x != 0; ++x) {
        inter[x + width - 1 + y * stride] = win.GetMaximum();
        if (x + width < nx) win.AddValue(value_at(x + width, y));
      }
      for (int x = 0; x < nx - width; ++x) {
        inter[x + width - 1 + y * stride] = win.GetMaximum();
        win.RemoveValue(value_at(x, y));
        win.AddValue(value_at(x + width, y));
      }
      for (int x = std::max(nx - width, 0); x != nx; ++x) {
        inter[x + width - 1 + y * stride] = win.GetMaximum();
        win.RemoveValue(value_at(x, y));
      }
      ORACLE_CHECK(win.empty());
    }
    // Column pass + quantisation.
    for (int x = 0; x != wide_x_; ++x) {
      SlidingWindowMaximum win;
      win.AddValue(inter[x]);
      for (int y = -width + 1; y != 0; ++y) {
        cells_[x + (y + width - 1) * stride] = ComputeCellValue(win.GetMaximum());
        if (y + width < ny) win.AddValue(inter[x + (y + width) * stride]);
      }
      for (int y = 0; y < ny - width; ++y) {
        cells_[x + (y + width - 1) * stride] = ComputeCellValue(win.GetMaximum());
        win.RemoveValue(inter[x + y * stride]);
        win.AddValue(inter[x + (y + width) * stride]);
      }
      for (int y = std::max(ny - width, 0); y != ny; ++y) {
        cells_[x + (y + width - 1) * stride] = ComputeCellValue(win.GetMaximum());
        win.RemoveValue(inter[x + y * stride]);
      }
      ORACLE_CHECK(win.empty());
    }
  }

  // fast_correlative_scan_matcher_2d.h:56-71
  int GetValue(const Array2i& xy) const {
    const int lx = xy.x - offset_x_, ly = xy.y - offset_y_;
    if (static_cast<unsigned>(lx) >= static_cast<unsigned>(wide_x_) ||
        static_cast<unsigned>(ly) >= static_cast<unsigned>(wide_y_)) {
      return 0;
    }
    return cells_[lx + ly * wide_x_];
  }
  // fast_correlative_scan_matcher_2d.h:74-76
  float ToScore(const float value) const {
    return min_score_ + value * ((max_score_ - min_score_) / 255.f);
  }
  int wide_x() const { return wide_x_; }
  int wide_y() const { return wide_y_; }
  const std::vector<uint8_t>& cells() const { return cells_; }

 private:
  // fast_correlative_scan_matcher_2d.cc:163-169
  uint8_t ComputeCellValue(const float probability) const {
    const int v = RoundToInt((probability - min_score_) * (255.f / (max_score_ - min_score_)));
    ORACLE_CHECK(v >= 0 && v <= 255);
    return static_cast<uint8_t>(v);
  }
  int offset_x_, offset_y_;
  int wide_x_, wide_y_;
  float min_score_, max_score_;
  std::vector<uint8_t> cells_;
};

// fast_correlative_scan_matcher_2d.cc:171-186
class PrecomputationGridStack2D {
 public:
  template <typename Grid>
  PrecomputationGridStack2D(const Grid& grid, const int branch_and_bound_depth) {
    ORACLE_CHECK(branch_and_bound_depth >= 1);
    std::vector<float> scratch;
    const CellLimits limits = grid.limits().cell_limits;
    grids_.reserve(branch_and_bound_depth);
    for (int i = 0; i != branch_and_bound_depth; ++i)
      grids_.emplace_back(grid, limits, 1 << i, &scratch);
  }
  const PrecomputationGrid2D& Get(int i) const { return grids_[i]; }
  int max_depth() const { return static_cast<int>(grids_.size()) - 1; }

 private:
  std::vector<PrecomputationGrid2D> grids_;
};

struct FastCorrelativeScanMatcherOptions2D {
  double linear_search_window = 7.;
  double angular_search_window = 30. * M_PI / 180.;
  int branch_and_bound_depth = 7;
};

// Work counters (not in the reference): candidate scores evaluated per level, for
// the candidate-scores/s metric of BASELINE.md.
struct MatchStats {
  uint64_t scored_top = 0;     // lowest-resolution candidates scored
  uint64_t scored_lower = 0;   // candidates scored inside BranchAndBound
  uint64_t num_scans = 0;
  uint64_t num_points = 0;
  uint64_t scored_at_depth[16] = {0};  // candidates scored with the grid of that depth
};

class FastCorrelativeScanMatcher2D {
 public:
  template <typename Grid>
  FastCorrelativeScanMatcher2D(const Grid& grid,
                               const FastCorrelativeScanMatcherOptions2D& options)
      : options_(options), limits_(grid.limits()), stack_(grid, options.branch_and_bound_depth) {}

  const PrecomputationGridStack2D& stack() const { return stack_; }

  struct Winner {  // extra observability for bit-level parity checks
    bool found = false;
    int scan_index = 0, x_index_offset = 0, y_index_offset = 0;
    float score = 0.f;
  };

  // fast_correlative_scan_matcher_2d.cc:198-209
  bool Match(const Rigid2d& initial, const PointCloud& cloud, const float min_score, float* score,
             Rigid2d* pose, Winner* winner = nullptr, MatchStats* stats = nullptr) const {
    const SearchParameters sp(options_.linear_search_window, options_.angular_search_window, cloud,
                              limits_.resolution);
    return MatchWithSearchParameters(sp, initial, cloud, min_score, score, pose, winner, stats);
  }
  // fast_correlative_scan_matcher_2d.cc:211-225
  bool MatchFullSubmap(const PointCloud& cloud, const float min_score, float* score, Rigid2d* pose,
                       Winner* winner = nullptr, MatchStats* stats = nullptr) const {
    const SearchParameters sp(1e6 * limits_.resolution, M_PI, cloud, limits_.resolution);
    Rigid2d center;
    center.tx = limits_.max_x - (0.5 * limits_.resolution) * limits_.cell_limits.num_y_cells;
    center.ty = limits_.max_y - (0.5 * limits_.resolution) * limits_.cell_limits.num_x_cells;
    center.angle = 0.;
    return MatchWithSearchParameters(sp, center, cloud, min_score, score, pose, winner, stats);
  }

 private:
  // fast_correlative_scan_matcher_2d.cc:227-262
  bool MatchWithSearchParameters(SearchParameters sp, const Rigid2d& initial,
                                 const PointCloud& cloud, const float min_score, float* score,
                                 Rigid2d* pose, Winner* winner, MatchStats* stats) const {
    ORACLE_CHECK(score != nullptr && pose != nullptr);
    const double initial_rotation = initial.angle;
    const PointCloud rotated = RotatePointCloudYaw(cloud, static_cast<float>(initial_rotation));
    const std::vector<PointCloud> rotated_scans = GenerateRotatedScans(rotated, sp);
    const std::vector<DiscreteScan2D> discrete = DiscretizeScans(
        limits_, rotated_scans, static_cast<float>(initial.tx), static_cast<float>(initial.ty));
    sp.ShrinkToFit(discrete, limits_.cell_limits);
    if (stats) {
      stats->num_scans += sp.num_scans;
      stats->num_points += cloud.size();
    }
    const std::vector<Candidate2D> lowest = ComputeLowestResolutionCandidates(discrete, sp, stats);
    const Candidate2D best = BranchAndBound(discrete, sp, lowest, stack_.max_depth(), min_score, stats);
    if (winner) {
      winner->found = best.score > min_score;
      winner->scan_index = best.scan_index;
      winner->x_index_offset = best.x_index_offset;
      winner->y_index_offset = best.y_index_offset;
      winner->score = best.score;
    }
    if (best.score > min_score) {
      *score = best.score;
      pose->tx = initial.tx + best.x;
      pose->ty = initial.ty + best.y;
      pose->angle = initial_rotation + best.orientation;  // Rotation2Dd * Rotation2Dd
      return true;
    }
    return false;
  }

  // fast_correlative_scan_matcher_2d.cc:264-312
  std::vector<Candidate2D> ComputeLowestResolutionCandidates(
      const std::vector<DiscreteScan2D>& discrete, const SearchParameters& sp,
      MatchStats* stats) const {
    const int step = 1 << stack_.max_depth();
    size_t count = 0;
    for (int s = 0; s != sp.num_scans; ++s) {
      const LinearBounds& lb = sp.linear_bounds[s];
      const int nx = (lb.max_x - lb.min_x + step) / step;
      const int ny = (lb.max_y - lb.min_y + step) / step;
      count += static_cast<size_t>(nx * ny);
    }
    std::vector<Candidate2D> cands;
    cands.reserve(count);
    for (int s = 0; s != sp.num_scans; ++s) {
      const LinearBounds& lb = sp.linear_bounds[s];
      for (int xo = lb.min_x; xo <= lb.max_x; xo += step)
        for (int yo = lb.min_y; yo <= lb.max_y; yo += step) cands.emplace_back(s, xo, yo, sp);
    }
    ORACLE_CHECK(cands.size() == count);
    if (stats) {
      stats->scored_top += cands.size();
      stats->scored_at_depth[stack_.max_depth() & 15] += cands.size();
    }
    ScoreCandidates(stack_.Get(stack_.max_depth()), discrete, &cands);
    return cands;
  }

  // fast_correlative_scan_matcher_2d.cc:314-333.  The trailing std::sort is the
  // literal, non-stable libstdc++ introsort of the reference.
  void ScoreCandidates(const PrecomputationGrid2D& pg, const std::vector<DiscreteScan2D>& discrete,
                       std::vector<Candidate2D>* cands) const {
    for (Candidate2D& c : *cands) {
      int sum = 0;
      for (const Array2i& xy : discrete[c.scan_index]) {
        Array2i p;
        p.x = xy.x + c.x_index_offset;
        p.y = xy.y + c.y_index_offset;
        sum += pg.GetValue(p);
      }
      c.score = pg.ToScore(sum / static_cast<float>(discrete[c.scan_index].size()));
    }
    std::sort(cands->begin(), cands->end(), std::greater<Candidate2D>());
  }

  // fast_correlative_scan_matcher_2d.cc:335-378
  Candidate2D BranchAndBound(const std::vector<DiscreteScan2D>& discrete, const SearchParameters& sp,
                             const std::vector<Candidate2D>& cands, const int depth,
                             float min_score, MatchStats* stats) const {
    if (depth == 0) return *cands.begin();
    Candidate2D best(0, 0, 0, sp);
    best.score = min_score;
    for (const Candidate2D& c : cands) {
      if (c.score <= min_score) break;
      std::vector<Candidate2D> children;
      const int half = 1 << (depth - 1);
      for (int dx : {0, half}) {
        if (c.x_index_offset + dx > sp.linear_bounds[c.scan_index].max_x) break;
        for (int dy : {0, half}) {
          if (c.y_index_offset + dy > sp.linear_bounds[c.scan_index].max_y) break;
          children.emplace_back(c.scan_index, c.x_index_offset + dx, c.y_index_offset + dy, sp);
        }
      }
      if (stats) {
        stats->scored_lower += children.size();
        stats->scored_at_depth[(depth - 1) & 15] += children.size();
      }
      ScoreCandidates(stack_.Get(depth - 1), discrete, &children);
      best = std::max(best, BranchAndBound(discrete, sp, children, depth - 1, best.score, stats));
    }
    return best;
  }

  FastCorrelativeScanMatcherOptions2D options_;
  MapLimits limits_;
  PrecomputationGridStack2D stack_;
};

// ---------------------------------------------------------------------------
// Real-time (exhaustive) matcher.  real_time_correlative_scan_matcher_2d.{h,cc}
// ---------------------------------------------------------------------------
struct RealTimeCorrelativeScanMatcherOptions {
  double linear_search_window = 0.1;
  double angular_search_window = 20. * M_PI / 180.;
  double translation_delta_cost_weight = 1e-1;
  double rotation_delta_cost_weight = 1e-1;
};

class RealTimeCorrelativeScanMatcher2D {
 public:
  explicit RealTimeCorrelativeScanMatcher2D(const RealTimeCorrelativeScanMatcherOptions& o)
      : options_(o) {}

  // real_time_correlative_scan_matcher_2d.cc:83-115
  std::vector<Candidate2D> GenerateExhaustiveSearchCandidates(const SearchParameters& sp) const {
    std::vector<Candidate2D> cands;
    for (int s = 0; s != sp.num_scans; ++s) {
      const LinearBounds& lb = sp.linear_bounds[s];
      for (int xo = lb.min_x; xo <= lb.max_x; ++xo)
        for (int yo = lb.min_y; yo <= lb.max_y; ++yo) cands.emplace_back(s, xo, yo, sp);
    }
    return cands;
  }

  // real_time_correlative_scan_matcher_2d.cc:61-75 (probability grid)
  static float ComputeCandidateScore(const ProbabilityGrid& grid, const DiscreteScan2D& scan,
                                     const int x_index_offset, const int y_index_offset) {
    float s = 0.f;
    for (const Array2i& xy : scan) {
      Array2i p;
      p.x = xy.x + x_index_offset;
      p.y = xy.y + y_index_offset;
      s += grid.GetProbability(p);
    }
    s /= static_cast<float>(scan.size());
    ORACLE_CHECK(s > 0.f);
    return s;
  }
  // real_time_correlative_scan_matcher_2d.cc:38-59 (TSDF)
  static float ComputeCandidateScore(const TSDF2D& tsdf, const DiscreteScan2D& scan,
                                     const int x_index_offset, const int y_index_offset) {
    float candidate_score = 0.f;
    float summed_weight = 0.f;
    for (const Array2i& xy : scan) {
      Array2i p;
      p.x = xy.x + x_index_offset;
      p.y = xy.y + y_index_offset;
      const std::pair<float, float> tsd_and_weight = tsdf.GetTSDAndWeight(p);
      const float normalized_tsd_score =
          (tsdf.GetMaxCorrespondenceCost() - std::abs(tsd_and_weight.first)) /
          tsdf.GetMaxCorrespondenceCost();
      const float weight = tsd_and_weight.second;
      candidate_score += normalized_tsd_score * weight;
      summed_weight += weight;
    }
    if (summed_weight == 0.f) return 0.f;
    candidate_score /= summed_weight;
    ORACLE_CHECK(candidate_score >= 0.f);
    return candidate_score;
  }

  // real_time_correlative_scan_matcher_2d.cc:151-176
  template <typename Grid>
  void ScoreCandidates(const Grid& grid, const std::vector<DiscreteScan2D>& discrete,
                       const SearchParameters& sp, std::vector<Candidate2D>* cands) const {
    (void)sp;
    for (Candidate2D& c : *cands) {
      c.score = ComputeCandidateScore(grid, discrete[c.scan_index], c.x_index_offset, c.y_index_offset);
      const double t = std::hypot(c.x, c.y) * options_.translation_delta_cost_weight +
                       std::abs(c.orientation) * options_.rotation_delta_cost_weight;
      c.score *= std::exp(-(t * t));  // float *= double
    }
  }

  // real_time_correlative_scan_matcher_2d.cc:117-149
  template <typename Grid>
  double Match(const Rigid2d& initial, const PointCloud& cloud, const Grid& grid,
               Rigid2d* pose, FastCorrelativeScanMatcher2D::Winner* winner = nullptr) const {
    ORACLE_CHECK(pose != nullptr);
    const double initial_rotation = initial.angle;
    const PointCloud rotated = RotatePointCloudYaw(cloud, static_cast<float>(initial_rotation));
    const SearchParameters sp(options_.linear_search_window, options_.angular_search_window,
                              rotated, grid.limits().resolution);
    const std::vector<PointCloud> scans = GenerateRotatedScans(rotated, sp);
    const std::vector<DiscreteScan2D> discrete = DiscretizeScans(
        grid.limits(), scans, static_cast<float>(initial.tx), static_cast<float>(initial.ty));
    std::vector<Candidate2D> cands = GenerateExhaustiveSearchCandidates(sp);
    ScoreCandidates(grid, discrete, sp, &cands);
    const Candidate2D& best = *std::max_element(cands.begin(), cands.end());
    pose->tx = initial.tx + best.x;
    pose->ty = initial.ty + best.y;
    pose->angle = initial_rotation + best.orientation;
    if (winner) {
      winner->found = true;
      winner->scan_index = best.scan_index;
      winner->x_index_offset = best.x_index_offset;
      winner->y_index_offset = best.y_index_offset;
      winner->score = best.score;
    }
    return best.score;
  }

 private:
  RealTimeCorrelativeScanMatcherOptions options_;
};

}  // namespace oracle

#endif  // ORACLE_CSM_ORACLE_H_
